// jsonl.cu -- index builder, host half.  Pure host C++ (no kernels; compiled with the library).
//
// Reads the JSONL files src/encode.py:353-359,426 writes -- one object per line, {"id": ..., "content": "", "vector":
// {term: int, ...}} -- the input of `pyserini.index.lucene --collection JsonVectorCollection --impact --pretokenized`
// (reference scripts/sparse_index.sh:12-18), and produces what the Python builder (index.py, ImpactCollection
// .from_vectors) produces, with the same numbering:
//   S1  tf(t,d) = vector[t]; non-positive weights are never indexed; a key holding whitespace contributes each of its
//       whitespace-separated pieces (Anserini repeats the key `value` times and whitespace-analyses the pseudo-document);
//       duplicate keys of one JSON object: the last one wins (json.loads);
//   S2  terms are taken verbatim;
//   S8  internal doc id = rank of the external docid string in UTF-8 byte order;
//   term ids are assigned in order of first appearance (document order, then key order inside the document).
// Output is doc-major (file order) for msr_invert_docs, plus the two string tables.
#include <errno.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <string>
#include <unordered_map>
#include <vector>

#include "common.cuh"

struct msr_collection {
    std::vector<std::string> docids;          // file order
    std::vector<uint32_t> doc_rank;           // file order -> internal id (S8)
    std::vector<std::string> terms;           // term id -> string
    std::vector<uint64_t> doc_indptr;         // [n_docs + 1]
    std::vector<uint32_t> post_terms;
    std::vector<uint16_t> post_impacts;
};

namespace msr {
namespace {

struct Parser {
    const char* p;
    const char* end;
    const char* file;
    uint64_t line;
    std::string err;

    bool fail(const char* what) {
        char buf[512];
        snprintf(buf, sizeof(buf), "%s:%llu: %s", file, (unsigned long long)line, what);
        err = buf;
        return false;
    }
    void ws() { while (p < end && (*p == ' ' || *p == '\t' || *p == '\r' || *p == '\n')) ++p; }
    bool lit(char c) { ws(); if (p < end && *p == c) { ++p; return true; } return false; }

    static void put_utf8(std::string& o, uint32_t c) {
        if (c < 0x80) o.push_back((char)c);
        else if (c < 0x800) { o.push_back((char)(0xC0 | (c >> 6))); o.push_back((char)(0x80 | (c & 0x3F))); }
        else if (c < 0x10000) { o.push_back((char)(0xE0 | (c >> 12))); o.push_back((char)(0x80 | ((c >> 6) & 0x3F))); o.push_back((char)(0x80 | (c & 0x3F))); }
        else { o.push_back((char)(0xF0 | (c >> 18))); o.push_back((char)(0x80 | ((c >> 12) & 0x3F))); o.push_back((char)(0x80 | ((c >> 6) & 0x3F))); o.push_back((char)(0x80 | (c & 0x3F))); }
    }
    bool hex4(uint32_t* out) {
        if (end - p < 4) return fail("truncated \\u escape");
        uint32_t v = 0;
        for (int i = 0; i < 4; ++i) {
            const char c = p[i];
            v <<= 4;
            if (c >= '0' && c <= '9') v |= c - '0';
            else if (c >= 'a' && c <= 'f') v |= c - 'a' + 10;
            else if (c >= 'A' && c <= 'F') v |= c - 'A' + 10;
            else return fail("bad \\u escape");
        }
        p += 4;
        *out = v;
        return true;
    }
    bool string(std::string& o) {
        o.clear();
        ws();
        if (p >= end || *p != '"') return fail("expected a string");
        ++p;
        while (p < end) {
            const char c = *p++;
            if (c == '"') return true;
            if (c != '\\') { o.push_back(c); continue; }
            if (p >= end) break;
            const char e = *p++;
            switch (e) {
                case '"': o.push_back('"'); break;
                case '\\': o.push_back('\\'); break;
                case '/': o.push_back('/'); break;
                case 'b': o.push_back('\b'); break;
                case 'f': o.push_back('\f'); break;
                case 'n': o.push_back('\n'); break;
                case 'r': o.push_back('\r'); break;
                case 't': o.push_back('\t'); break;
                case 'u': {
                    uint32_t c1 = 0;
                    if (!hex4(&c1)) return false;
                    if (c1 >= 0xD800 && c1 <= 0xDBFF) {
                        uint32_t c2 = 0;
                        if (end - p >= 6 && p[0] == '\\' && p[1] == 'u') { p += 2; if (!hex4(&c2)) return false; }
                        if (c2 < 0xDC00 || c2 > 0xDFFF) return fail("lone surrogate in a string (not encodable as UTF-8)");
                        c1 = 0x10000 + ((c1 - 0xD800) << 10) + (c2 - 0xDC00);
                    } else if (c1 >= 0xDC00 && c1 <= 0xDFFF) {
                        return fail("lone surrogate in a string (not encodable as UTF-8)");
                    }
                    put_utf8(o, c1);
                    break;
                }
                default: return fail("bad escape in a string");
            }
        }
        return fail("unterminated string");
    }
    // JSON number -> what Python's int(v) gives for the parsed value (ints exactly, floats truncated toward zero)
    bool number(long long* out) {
        ws();
        const char* s = p;
        if (p < end && (*p == '-' || *p == '+')) ++p;
        bool is_float = false;
        while (p < end && ((*p >= '0' && *p <= '9') || *p == '.' || *p == 'e' || *p == 'E' || *p == '-' || *p == '+')) {
            if (*p == '.' || *p == 'e' || *p == 'E') is_float = true;
            ++p;
        }
        if (p == s) return fail("expected a number");
        const std::string tok(s, p - s);
        if (is_float) {
            const double d = strtod(tok.c_str(), nullptr);
            if (!(d > -9.2e18 && d < 9.2e18)) return fail("weight out of range");
            *out = (long long)d;                                 // truncation toward zero, as int(float)
        } else {
            errno = 0;
            *out = strtoll(tok.c_str(), nullptr, 10);
            if (errno) return fail("weight out of range");
        }
        return true;
    }
    bool skip_value() {
        ws();
        if (p >= end) return fail("unexpected end of line");
        std::string tmp;
        if (*p == '"') return string(tmp);
        if (*p == '{' || *p == '[') {
            const char open = *p, close = open == '{' ? '}' : ']';
            ++p;
            if (lit(close)) return true;
            for (;;) {
                if (open == '{') { if (!string(tmp)) return false; if (!lit(':')) return fail("expected ':'"); }
                if (!skip_value()) return false;
                if (lit(',')) continue;
                if (lit(close)) return true;
                return fail("expected ',' or a closing bracket");
            }
        }
        const char* s = p;
        while (p < end && *p != ',' && *p != '}' && *p != ']' && *p != ' ' && *p != '\t' && *p != '\r' && *p != '\n') ++p;
        return p > s ? true : fail("unexpected character");
    }
};

// Python's str.split() separators, on a UTF-8 string: returns the byte length of the whitespace code point at s (0 = none)
inline int py_space(const unsigned char* s, const unsigned char* end) {
    const unsigned c = s[0];
    if (c < 0x80) return ((c >= 0x09 && c <= 0x0D) || (c >= 0x1C && c <= 0x20)) ? 1 : 0;
    if (c == 0xC2 && end - s >= 2) return (s[1] == 0x85 || s[1] == 0xA0) ? 2 : 0;
    if (end - s >= 3) {
        const unsigned cp = ((c & 0x0F) << 12) | ((s[1] & 0x3F) << 6) | (s[2] & 0x3F);
        if ((c & 0xF0) == 0xE0 && (cp == 0x1680 || (cp >= 0x2000 && cp <= 0x200A) || cp == 0x2028 || cp == 0x2029 || cp == 0x202F ||
                                   cp == 0x205F || cp == 0x3000)) return 3;
    }
    return 0;
}

}  // namespace
}  // namespace msr

using namespace msr;

namespace {

struct KV { std::string key; long long val; };

// one JSONL line -> (id, vector items in dict order with last-wins values); returns false with P.err set
bool parse_line(Parser& P, std::string& id, bool* has_id, std::vector<KV>& items) {
    items.clear();
    *has_id = false;
    if (!P.lit('{')) return P.fail("expected a JSON object");
    std::string key, sval;
    if (!P.lit('}')) {
        for (;;) {
            if (!P.string(key)) return false;
            if (!P.lit(':')) return P.fail("expected ':'");
            if (key == "id") {
                P.ws();
                if (P.p < P.end && *P.p == '"') {
                    if (!P.string(id)) return false;
                } else {
                    const char* s = P.p;
                    long long v = 0;
                    if (!P.number(&v)) return false;
                    for (const char* c = s; c < P.p; ++c)
                        if (*c == '.' || *c == 'e' || *c == 'E') return P.fail("non-integer numeric id is not supported");
                    id = std::to_string(v);                       // str(int)
                }
                *has_id = true;
            } else if (key == "vector") {
                items.clear();                                   // a repeated "vector" key: the last object wins
                P.ws();
                if (P.p < P.end && *P.p == '{') {
                    ++P.p;
                    if (!P.lit('}')) {
                        for (;;) {
                            KV kv;
                            if (!P.string(kv.key)) return false;
                            if (!P.lit(':')) return P.fail("expected ':'");
                            P.ws();
                            if (P.p < P.end && *P.p == '"') {         // int("12") also works in the Python builder
                                if (!P.string(sval)) return false;
                                char* e = nullptr;
                                kv.val = strtoll(sval.c_str(), &e, 10);
                                if (sval.empty() || *e) return P.fail("vector weight is not an integer");
                            } else if (!P.number(&kv.val)) {
                                return false;
                            }
                            bool dup = false;
                            for (KV& o : items)
                                if (o.key == kv.key) { o.val = kv.val; dup = true; break; }   // dict: position kept, value replaced
                            if (!dup) items.push_back(std::move(kv));
                            if (P.lit(',')) continue;
                            if (P.lit('}')) break;
                            return P.fail("expected ',' or '}' in the vector");
                        }
                    }
                } else {
                    const char* s = P.p;
                    if (!P.skip_value()) return false;
                    if (!(P.p - s == 4 && !strncmp(s, "null", 4))) return P.fail("\"vector\" must be an object");
                }
            } else if (!P.skip_value()) {
                return false;
            }
            if (P.lit(',')) continue;
            if (P.lit('}')) break;
            return P.fail("expected ',' or '}'");
        }
    }
    P.ws();
    if (P.p != P.end) return P.fail("extra data after the JSON object");
    return true;
}

}  // namespace

extern "C" int msr_jsonl_parse(const char* const* paths, uint32_t n_paths, msr_collection** out) {
    MSR_REQUIRE(out, "msr_jsonl_parse: out is NULL");
    *out = nullptr;
    MSR_REQUIRE(paths && n_paths > 0, "msr_jsonl_parse: no input files");
    msr_collection* C = new (std::nothrow) msr_collection();
    if (!C) { set_error("msr_jsonl_parse: out of host memory"); return MSR_ERR_NOMEM; }
    std::unordered_map<std::string, uint32_t> term_id;
    std::vector<KV> items;
    std::vector<std::pair<std::string, long long>> tf;          // insertion-ordered Counter of one document
    std::unordered_map<std::string, uint32_t> tf_pos;
    std::string id, piece;
    C->doc_indptr.push_back(0);
    char* line = nullptr;
    size_t cap = 0;
    int rc = MSR_OK;
    for (uint32_t f = 0; f < n_paths && rc == MSR_OK; ++f) {
        FILE* fp = fopen(paths[f], "rb");
        if (!fp) { set_error("msr_jsonl_parse: cannot open %s", paths[f]); rc = MSR_ERR_INVALID; break; }
        uint64_t ln = 0;
        ssize_t n;
        while ((n = getline(&line, &cap, fp)) >= 0) {
            ++ln;
            Parser P{line, line + n, paths[f], ln, std::string()};
            P.ws();
            if (P.p == P.end) continue;                          // blank line
            bool has_id = false;
            if (!parse_line(P, id, &has_id, items)) { set_error("msr_jsonl_parse: %s", P.err.c_str()); rc = MSR_ERR_INVALID; break; }
            if (!has_id) { set_error("msr_jsonl_parse: %s:%llu: no \"id\" key", paths[f], (unsigned long long)ln); rc = MSR_ERR_INVALID; break; }
            tf.clear(); tf_pos.clear();
            for (const KV& kv : items) {
                if (kv.val <= 0) continue;                       // S1
                const unsigned char* s = (const unsigned char*)kv.key.data();
                const unsigned char* e = s + kv.key.size();
                while (s < e) {                                  // str.split()
                    int w;
                    while (s < e && (w = py_space(s, e)) > 0) s += w;
                    const unsigned char* b = s;
                    while (s < e && py_space(s, e) == 0) ++s;
                    if (s == b) break;
                    piece.assign((const char*)b, s - b);
                    auto it = tf_pos.find(piece);
                    if (it == tf_pos.end()) { tf_pos.emplace(piece, (uint32_t)tf.size()); tf.emplace_back(piece, kv.val); }
                    else tf[it->second].second += kv.val;
                }
            }
            for (const auto& tv : tf) {
                if (tv.second > 65535) {
                    set_error("msr_jsonl_parse: impact %lld of term '%s' in doc '%s' exceeds 65535", tv.second, tv.first.c_str(), id.c_str());
                    rc = MSR_ERR_RANGE;
                    break;
                }
                auto it = term_id.find(tv.first);
                uint32_t t;
                if (it == term_id.end()) { t = (uint32_t)C->terms.size(); term_id.emplace(tv.first, t); C->terms.push_back(tv.first); }
                else t = it->second;
                C->post_terms.push_back(t);
                C->post_impacts.push_back((uint16_t)tv.second);
            }
            if (rc) break;
            C->docids.push_back(id);
            C->doc_indptr.push_back(C->post_terms.size());
        }
        fclose(fp);
    }
    free(line);
    if (rc == MSR_OK) {
        // S8: internal id = rank of the docid string in UTF-8 byte order; duplicates are an error
        const size_t n = C->docids.size();
        std::vector<uint32_t> order(n);
        for (size_t i = 0; i < n; ++i) order[i] = (uint32_t)i;
        std::sort(order.begin(), order.end(), [&](uint32_t a, uint32_t b) { return C->docids[a] < C->docids[b]; });
        C->doc_rank.resize(n);
        for (size_t r = 0; r < n; ++r) {
            if (r && C->docids[order[r]] == C->docids[order[r - 1]]) {
                set_error("msr_jsonl_parse: duplicate document id '%s'", C->docids[order[r]].c_str());
                rc = MSR_ERR_INVALID;
                break;
            }
            C->doc_rank[order[r]] = (uint32_t)r;
        }
    }
    if (rc) { delete C; return rc; }
    *out = C;
    return MSR_OK;
}

extern "C" int msr_collection_destroy(msr_collection* c) {
    delete c;
    return MSR_OK;
}

// info: n_docs, n_terms, nnz, bytes of all docid strings, bytes of all term strings
extern "C" int msr_collection_info(const msr_collection* c, uint64_t info[5]) {
    MSR_REQUIRE(c && info, "msr_collection_info: NULL argument");
    uint64_t db = 0, tb = 0;
    for (const std::string& s : c->docids) db += s.size();
    for (const std::string& s : c->terms) tb += s.size();
    info[0] = c->docids.size(); info[1] = c->terms.size(); info[2] = c->post_terms.size(); info[3] = db; info[4] = tb;
    return MSR_OK;
}

extern "C" int msr_collection_export(const msr_collection* c, uint64_t* doc_indptr, uint32_t* terms, uint16_t* impacts,
                                     uint32_t* doc_rank, char* docid_blob, uint64_t* docid_off, char* term_blob,
                                     uint64_t* term_off) {
    MSR_REQUIRE(c, "msr_collection_export: NULL handle");
    if (doc_indptr) memcpy(doc_indptr, c->doc_indptr.data(), c->doc_indptr.size() * 8);
    if (terms && !c->post_terms.empty()) memcpy(terms, c->post_terms.data(), c->post_terms.size() * 4);
    if (impacts && !c->post_impacts.empty()) memcpy(impacts, c->post_impacts.data(), c->post_impacts.size() * 2);
    if (doc_rank && !c->doc_rank.empty()) memcpy(doc_rank, c->doc_rank.data(), c->doc_rank.size() * 4);
    auto blob = [](const std::vector<std::string>& v, char* b, uint64_t* off) {
        uint64_t o = 0;
        for (size_t i = 0; i < v.size(); ++i) {
            if (off) off[i] = o;
            if (b) memcpy(b + o, v[i].data(), v[i].size());
            o += v[i].size();
        }
        if (off) off[v.size()] = o;
    };
    if (docid_blob || docid_off) blob(c->docids, docid_blob, docid_off);
    if (term_blob || term_off) blob(c->terms, term_blob, term_off);
    return MSR_OK;
}
