// sparse_dense_mma.cuh -- dense phase of the sparse scorer on the integer tensor pipe (included by sparse_kernel.cuh).
//
// The dense phase adds impact * weight for every (doc of the tile, dense term of the query).  As dp4a with a one-byte
// selector it is ONE fma-pipe instruction per pair, and the fma pipe issues a warp instruction every 2 cycles per
// scheduler (csrc/microbench.cu: 1.98 dp4a / clock / SM) -- half of the kernel's samples sat there.  Here the same sums
// come from mma.sync.m16n8k32 (u8 x u8 -> s32; SASS IMMA.16832.U8.U8, measured 0.38 / clock / SM = 197 pairs / clock /
// SM used this way against 63 for dp4a):
//
//   * four dense terms of the query form a QUAD; a quad's four u8 weights are one 32-bit word W;
//   * A[m][k] (16 x 32, row-major) is cut into 8 column groups of 4: A[m][4n .. 4n+3] = the quad's four impacts of ONE doc,
//     so the 16 x 8 (m, n) cells are 128 docs; B[k][n] = W's byte k % 4 when k / 4 == n, else 0 (block diagonal), hence
//     D[m][n] += sum over the quad's terms of impact(doc(m, n), term) * weight(term): 128 docs x 4 terms per instruction;
//   * the impact rows are term-major (one u8 per doc), so a lane loads 8 docs of one term with an 8-byte load, as before,
//     and a 4 x 4 byte transpose (8 PRMT, alu pipe -- idle so far) turns the four terms' words into four doc words;
//   * lane L = 4g + t supplies docs 8L .. 8L+7 of the warp's 256-doc block (two mma tiles of 128) and receives
//     D[g][2t], D[g][2t+1], D[g+8][2t], D[g+8][2t+1], which are OTHER docs.  Instead of shuffling them back, the shared
//     memory accumulators are kept in that order: lane L owns accumulator POSITIONS 8L .. 8L+7 of the block, and
//     position <-> doc is a permutation of bits 1, 3, 4 of the doc's offset inside its 32-doc group (perm32 / unperm32).
//     The posting buckets carry positions instead of doc offsets (msr_sparse_index_build applies perm32 once), so the
//     scatter phase is unchanged, and the selection passes turn a position back into a doc id where they emit one.
//
// Weights of 256 and more and S6 terms (u16 rows) are added on the ALU as before, read in position order.
#pragma once

namespace msr {

// MEASURED (profiles/r2_dense_mma_experiment.json, same box, config 4): bit-exact, and SLOWER than the dp4a form -- scorer
// 19.95 ms against 19.25 ms (490 K against 508 K q/s).  The dense phase is not bound by the fma pipe (26 % busy over the
// kernel) but by L2 latency at 32 warps per SM; taking instructions out of it does not shorten the wait for the row
// loads.  The experiment stays in the tree behind -DMSR_DENSE_MMA=1 (build.build_variant); the default is the dp4a form.
#ifndef MSR_DENSE_MMA
#define MSR_DENSE_MMA 0
#endif

// doc offset inside the tile -> accumulator position (and back): bits 1 -> 4, 4 -> 3, 3 -> 1 of the offset
__host__ __device__ __forceinline__ uint32_t perm32(uint32_t x) {
#if MSR_DENSE_MMA
    return (x & ~0x1Au) | ((x & 0x02u) << 3) | ((x & 0x10u) >> 1) | ((x & 0x08u) >> 2);
#else
    return x;
#endif
}
__host__ __device__ __forceinline__ uint32_t unperm32(uint32_t p) {
#if MSR_DENSE_MMA
    return (p & ~0x1Au) | ((p & 0x10u) >> 3) | ((p & 0x08u) << 1) | ((p & 0x02u) << 2);
#else
    return p;
#endif
}

#ifdef __CUDACC__
// streaming loads of the index (read once per CTA: no L1 allocation)
__device__ __forceinline__ uint4 ld_stream_v4(const void* p) {
    uint4 v;
    asm("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
    return v;
}
__device__ __forceinline__ uint2 ld_stream_v2(const void* p) {
    uint2 v;
    asm("ld.global.nc.L1::no_allocate.v2.u32 {%0,%1}, [%2];" : "=r"(v.x), "=r"(v.y) : "l"(p));
    return v;
}
#endif

#if MSR_DENSE_MMA && defined(__CUDACC__)

// acc[0..3] are positions 4T .. 4T+3 of one mma tile T: D[g][2t] = position 0, D[g+8][2t] = 1, D[g][2t+1] = 2, D[g+8][2t+1] = 3
__device__ __forceinline__ void imma_quad(uint32_t (&acc)[8], const int T, uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3,
                                          uint32_t b0, uint32_t b1) {
    asm("mma.sync.aligned.m16n8k32.row.col.s32.u8.u8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+r"(acc[4 * T + 0]), "+r"(acc[4 * T + 2]), "+r"(acc[4 * T + 1]), "+r"(acc[4 * T + 3])
        : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

// N words (N terms x 4 docs, byte i of word j = impact of doc i for term j) -> 4 words (doc i: byte j = term j).
// Bytes of absent terms (N < 4) are arbitrary: their weight byte in B is 0.
template <int N>
__device__ __forceinline__ void quad_transpose(uint32_t w0, uint32_t w1, uint32_t w2, uint32_t w3, uint32_t& a0, uint32_t& a1,
                                               uint32_t& a2, uint32_t& a3) {
    if (N == 1) {
        a0 = w0; a1 = w0 >> 8; a2 = w0 >> 16; a3 = w0 >> 24;
    } else if (N == 2) {
        a0 = __byte_perm(w0, w1, 0x0040); a1 = __byte_perm(w0, w1, 0x0051);
        a2 = __byte_perm(w0, w1, 0x0062); a3 = __byte_perm(w0, w1, 0x0073);
    } else {
        const uint32_t t01 = __byte_perm(w0, w1, 0x5140);       // w0.b0 w1.b0 w0.b1 w1.b1
        const uint32_t t23 = __byte_perm(w0, w1, 0x7362);       // w0.b2 w1.b2 w0.b3 w1.b3
        if (N == 3) {
            a0 = __byte_perm(t01, w2, 0x0410); a1 = __byte_perm(t01, w2, 0x0532);
            a2 = __byte_perm(t23, w2, 0x0610); a3 = __byte_perm(t23, w2, 0x0732);
        } else {
            const uint32_t u01 = __byte_perm(w2, w3, 0x5140), u23 = __byte_perm(w2, w3, 0x7362);
            a0 = __byte_perm(t01, u01, 0x5410); a1 = __byte_perm(t01, u01, 0x7632);
            a2 = __byte_perm(t23, u23, 0x5410); a3 = __byte_perm(t23, u23, 0x7632);
        }
    }
}

// one quad of N terms over the thread's two slots (8 docs each = two mma tiles each); r = the terms' row offsets (bytes)
template <int N>
__device__ __forceinline__ void dense_quad(uint32_t (&pa)[8], uint32_t (&pb)[8], const char* r8, const uint32_t kB8, const uint4 r,
                                           const uint32_t b0, const uint32_t b1) {
    uint2 v0, v1 = make_uint2(0, 0), v2 = make_uint2(0, 0), v3 = make_uint2(0, 0);
    uint2 u0, u1 = make_uint2(0, 0), u2 = make_uint2(0, 0), u3 = make_uint2(0, 0);
    v0 = ld_stream_v2(r8 + r.x); u0 = ld_stream_v2(r8 + r.x + kB8);
    if (N > 1) { v1 = ld_stream_v2(r8 + r.y); u1 = ld_stream_v2(r8 + r.y + kB8); }
    if (N > 2) { v2 = ld_stream_v2(r8 + r.z); u2 = ld_stream_v2(r8 + r.z + kB8); }
    if (N > 3) { v3 = ld_stream_v2(r8 + r.w); u3 = ld_stream_v2(r8 + r.w + kB8); }
    uint32_t a0, a1, a2, a3;
    quad_transpose<N>(v0.x, v1.x, v2.x, v3.x, a0, a1, a2, a3); imma_quad(pa, 0, a0, a1, a2, a3, b0, b1);
    quad_transpose<N>(v0.y, v1.y, v2.y, v3.y, a0, a1, a2, a3); imma_quad(pa, 1, a0, a1, a2, a3, b0, b1);
    quad_transpose<N>(u0.x, u1.x, u2.x, u3.x, a0, a1, a2, a3); imma_quad(pb, 0, a0, a1, a2, a3, b0, b1);
    quad_transpose<N>(u0.y, u1.y, u2.y, u3.y, a0, a1, a2, a3); imma_quad(pb, 1, a0, a1, a2, a3, b0, b1);
}

// a term on the u16 rows (weight >= 256 or S6), read in position order: the thread's 8 positions are docs
// 32g + 16(t&1) + 2(t>>1) + {0,1} + 4T + 8*pair of its 32-doc group, i.e. every other 32-bit word of a 32-byte row chunk
__device__ __forceinline__ void dense_wide_term(uint32_t (&p)[8], const char* chunk, const uint32_t hi, const uint32_t wr) {
    const uint4 q0 = ld_stream_v4(chunk), q1 = ld_stream_v4(chunk + 16);
    // words (2 docs each): index 4*pair + 2*T + (t>>1)
    const uint32_t e[4] = {hi ? q0.y : q0.x, hi ? q0.w : q0.z, hi ? q1.y : q1.x, hi ? q1.w : q1.z};   // (T,pair) = (0,0) (1,0) (0,1) (1,1)
    const uint32_t w = wr & 0x7FFFFFFFu;
    const bool s6 = wr >> 31;
#pragma unroll
    for (int T = 0; T < 2; ++T)
#pragma unroll
        for (int pair = 0; pair < 2; ++pair) {
            const uint32_t x = e[2 * pair + T];
            const uint32_t lo = (x & 0xFFFFu) * w, up = (x >> 16) * w;
            p[4 * T + 2 * pair + 0] += s6 ? (uint32_t)__uint2float_rn(lo) : lo;
            p[4 * T + 2 * pair + 1] += s6 ? (uint32_t)__uint2float_rn(up) : up;
        }
}

#endif  // MSR_DENSE_MMA

}  // namespace msr
