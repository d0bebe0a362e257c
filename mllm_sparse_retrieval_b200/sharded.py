"""Doc-sharded exchange step: per-GPU top-k lists -> every GPU -> merge.

Two implementations: ``PeerExchange`` (default at N > 1 in bench.py) fuses the transfer into the scorer's tile-merge
epilogue as NVLink peer stores into symmetric memory, with no collective call on the path; ``PipelinedExchange`` is the
NCCL all-gather form, pipelined over query chunks, and the fallback when symmetric memory cannot be set up.

The document collection is split into contiguous internal-id ranges, one per GPU (the precedent in the reference is
FAISS ``IndexShards`` behind ``GpuMultipleClonerOptions.shard=True``, src/search.py:265-270).  Every rank scores the
whole query batch against its shard; the only exchange is an all-gather of ``[nq, k]`` (score, id) lists plus counts,
followed by ``msr_topk_merge`` with the same (score desc, id asc) key, so the result does not depend on the number
of shards.

``PipelinedExchange`` cuts the query batch into chunks: while the scorer works on chunk i+1 on the compute stream, the
all-gather and merge of chunk i run on a side stream (NCCL over NVLink needs a handful of SMs; the scorer keeps the
rest).  With one chunk it degenerates to score -> gather -> merge; with one rank it only runs the chunks.
Measured on config 4 (1 M docs, 10 K queries, k = 100): the exchange is ~4 % of the step at 2 GPUs and ~12 % at 8, and
smaller scorer launches lose about as much as the overlap wins (2 GPUs: 668 K QPS with 1 chunk, 640 K with 4; 8 GPUs:
1.85 M with 1, 1.84 M with 2), so the default is ONE chunk; dense search uses its natural query batches as chunks.
"""
from __future__ import annotations

import os

import torch
import torch.distributed as dist

from . import ops


class PipelinedExchange:
    def __init__(self, nq, k, device, world, *, chunks=1, min_chunk=1024, id_dtype=torch.int32, group=None):
        self.nq, self.k, self.device, self.world, self.group = nq, k, torch.device(device), world, group
        n_chunks = max(1, min(chunks, nq // max(1, min_chunk)))
        self.bounds = [nq * i // n_chunks for i in range(n_chunks + 1)]
        dev = self.device
        # local results of the whole batch (chunk views are contiguous row ranges)
        self.local = (torch.empty((nq, k), dtype=torch.float32, device=dev), torch.empty((nq, k), dtype=id_dtype, device=dev),
                      torch.empty((nq,), dtype=torch.int32, device=dev))
        self.final = self.local
        self.gathered = []
        if world > 1:
            self.final = (torch.empty((nq, k), dtype=torch.float32, device=dev), torch.empty((nq, k), dtype=torch.int32, device=dev),
                          torch.empty((nq,), dtype=torch.int32, device=dev))
            for lo, hi in self.chunks():
                n = hi - lo
                self.gathered.append((torch.empty((world, n, k), dtype=torch.float32, device=dev),
                                      torch.empty((world, n, k), dtype=torch.int32, device=dev),
                                      torch.empty((world, n), dtype=torch.int32, device=dev)))
            self.side = torch.cuda.Stream(device=dev)
            self.events = [torch.cuda.Event() for _ in self.gathered]

    def chunks(self):
        return list(zip(self.bounds[:-1], self.bounds[1:]))

    def local_views(self, lo, hi):
        return tuple(t[lo:hi] for t in self.local)

    def run(self, search_chunk):
        """search_chunk(i, lo, hi, out_views) scores queries [lo,hi) of the batch against the local shard on the
        current stream, writing (scores, ids, counts) into out_views.  Returns the merged (scores, ids, counts) of the
        whole batch, ready on the current stream."""
        if self.world == 1:
            for i, (lo, hi) in enumerate(self.chunks()):
                search_chunk(i, lo, hi, self.local_views(lo, hi))
            return self.local
        main = torch.cuda.current_stream(self.device)
        self.side.wait_stream(main)                              # buffers of the previous run are free again
        for i, (lo, hi) in enumerate(self.chunks()):
            ls, ld, lc = self.local_views(lo, hi)
            search_chunk(i, lo, hi, (ls, ld, lc))
            self.events[i].record(main)
            g_s, g_d, g_c = self.gathered[i]
            with torch.cuda.stream(self.side):
                self.side.wait_event(self.events[i])
                dist.all_gather_into_tensor(g_s, ls, group=self.group)
                dist.all_gather_into_tensor(g_d, ld if ld.dtype == torch.int32 else ld.to(torch.int32), group=self.group)
                dist.all_gather_into_tensor(g_c, lc, group=self.group)
                ops.topk_merge(g_s, g_d, g_c, self.k, out=tuple(t[lo:hi] for t in self.final))
        main.wait_stream(self.side)
        return self.final


def peer_layout(nq, k, world):
    """Symmetric buffer of ``PeerExchange`` in 4-byte units (every section starts on a 16-byte boundary):
    2 step buffers x world slots of [scores nq*k | docs nq*k | counts nq(+pad)], then 3 rotating threshold regions of
    [theta u64[nq] | quantiles u64[world][nq]], then world flags.  Returns (slot, theta_off, flag_off, total)."""
    al = lambda x: (x + 3) // 4 * 4
    slot = 2 * al(nq * k) + al(nq)
    theta_off = 2 * world * slot
    flag_off = theta_off + 3 * al(2 * nq * (1 + world))
    return slot, theta_off, flag_off, flag_off + al(world)


class PeerExchange:
    """The same exchange with the transfer FUSED into the scorer's last kernel: the tile-merge epilogue of every rank
    stores its [nq,k] result rows straight into that rank's slot of a symmetric-memory buffer on every other GPU (16-byte
    NVLink peer stores, ``msr_sparse_set_peer_outputs``); ONE more kernel (``msr_topk_merge_exchange``) then signals the
    peers, waits for their signals and merges the slots in place.  No NCCL call and no separate barrier launch is on the
    path.  Two step buffers alternate, so a rank that runs ahead writes the slots of step n+1 while a slower peer still
    merges step n; a rank can only reach step n+2 after every peer has signalled step n+1, i.e. finished merging step n.

    The scorer's pruning threshold is shared the same way (``msr_sparse_set_peer_theta``): theta[q] lives in the symmetric
    buffer and every raise is pushed to all peers, and so is every GPU's local ceil(k/world)-th best key per query -- the
    smallest of those over the GPUs bounds the collection's k-th key far better than any shard's own k-th key does -- so a
    shard with 6 tiles prunes like one with 41 and hands over short lists.  Three threshold
    regions rotate (step n uses n % 3 and clears (n + 2) % 3), so a peer that is one step ahead never has its pushes
    wiped, and no array is reused before every rank has finished with it.

    Raises (RuntimeError / ImportError) when symmetric memory cannot be set up; callers fall back to
    ``PipelinedExchange`` (NCCL)."""

    def __init__(self, nq, k, device, world, rank, group=None, share_theta=True):
        import torch.distributed._symmetric_memory as symm_mem
        self.nq, self.k, self.device, self.world, self.rank = nq, k, torch.device(device), world, rank
        group = group if group is not None else dist.group.WORLD
        self.slot, self.theta_off, self.flag_off, total = peer_layout(nq, k, world)
        self.buf = symm_mem.empty((total,), dtype=torch.int32, device=self.device)
        self.buf.zero_()
        self.hdl = symm_mem.rendezvous(self.buf, group)
        self.ptrs = [int(p) for p in self.hdl.buffer_ptrs]
        self.final = (torch.empty((nq, k), dtype=torch.float32, device=self.device),
                      torch.empty((nq, k), dtype=torch.int32, device=self.device),
                      torch.empty((nq,), dtype=torch.int32, device=self.device))
        self.step = 0
        self.share_theta = share_theta and os.environ.get("MSR_SPARSE_NO_PEER_THETA") != "1"   # developer knobs (A/B runs)
        self.share_quant = self.share_theta and os.environ.get("MSR_SPARSE_NO_PEER_QUANT") != "1"
        self.peers = [r for r in range(world) if r != rank]
        self.profile = os.environ.get("MSR_PX_PROFILE") == "1"         # CUDA events around the two halves of a step
        self._ev = []
        torch.cuda.synchronize(self.device)
        self.hdl.barrier(channel=0)                                     # every rank's buffer is zeroed before anyone writes

    def _slot_ptrs(self, base, parity, who):
        al = lambda x: (x + 3) // 4 * 4
        b = base + 4 * ((parity * self.world + who) * self.slot)
        return b, b + 4 * al(self.nq * self.k), b + 8 * al(self.nq * self.k)

    def _theta_ptr(self, base, which):
        """address of rotating threshold region ``which`` (theta[nq], then quant[world][nq] 8 * nq bytes further on)"""
        return base + 4 * (self.theta_off + which * ((2 * self.nq * (1 + self.world) + 3) // 4 * 4))

    def run(self, index, search, nq=None):
        """search(out_views) runs ONE device-path search of the whole batch on ``index`` (a ``SparseIndex``) writing
        (scores, ids, counts) into out_views; returns the merged result of all shards on the current stream.
        ``nq`` <= the capacity given at construction (default: the capacity) is the batch size of THIS call and must be
        the same on every rank; the slots keep their capacity-sized layout and only the first ``nq`` rows are used."""
        nq = self.nq if nq is None else int(nq)
        assert 0 < nq <= self.nq
        par = self.step & 1
        t_now, t_clear = self.step % 3, (self.step + 2) % 3
        self.step += 1
        al = lambda x: (x + 3) // 4 * 4
        peers = [self._slot_ptrs(self.ptrs[r], par, self.rank) for r in self.peers]
        index.set_peer_outputs([p[0] for p in peers], [p[1] for p in peers], [p[2] for p in peers])
        if self.share_theta:
            # the array two steps ahead is cleared now: nobody can be writing it yet, and nobody will read it before
            tw = al(2 * self.nq * (1 + self.world))
            o = self.theta_off + t_clear * tw
            self.buf[o:o + tw].zero_()
            mine = self._theta_ptr(self.ptrs[self.rank], t_now)
            theirs = [self._theta_ptr(self.ptrs[r], t_now) for r in self.peers]
            index.set_peer_theta(mine, theirs, self.nq, self.rank, self.world,
                                 local_quant=mine + 8 * self.nq if self.share_quant else 0, peer_quant=[t + 8 * self.nq for t in theirs])
        if self.profile:
            ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
            self._ev.append(ev)
            ev[0].record()
        try:
            # the local slot is the search's own output: views into the symmetric buffer
            o = self.slot * (par * self.world + self.rank)
            cap_nk = al(self.nq * self.k)
            out = (self.buf[o:o + nq * self.k].view(torch.float32).view(nq, self.k),
                   self.buf[o + cap_nk:o + cap_nk + nq * self.k].view(nq, self.k),
                   self.buf[o + 2 * cap_nk:o + 2 * cap_nk + nq])
            search(out)
        finally:
            index.set_peer_outputs()
            if self.share_theta:
                index.set_peer_theta()
        s0, d0, c0 = self._slot_ptrs(self.ptrs[self.rank], par, 0)
        final = self.final if nq == self.nq else tuple(t[:nq] for t in self.final)
        if self.profile:
            self._ev[-1][1].record()
        ops.topk_merge_exchange(s0, d0, c0, self.world, nq, self.k, self.slot, self.slot, self.k, final, self.device,
                                self.rank, self.ptrs[self.rank] + 4 * self.flag_off,
                                [self.ptrs[r] + 4 * self.flag_off for r in self.peers], self.peers, self.step)
        if self.profile:
            self._ev[-1][2].record()
        return final

    def exchange_rows(self, scores, docs, counts):
        """The exchange for a search that is already done (the dense flat-IP search): this rank's rows -- CUDA tensors
        scores f32 [nq,k], docs i32 [nq,k] (global ids), counts i32 [nq], sorted best-first -- are stored into its slot on
        every GPU (msr_exchange_publish), then one-warp flag kernel + merge (msr_topk_merge_exchange).  nq <= capacity, the
        same on every rank.  Returns the merged (scores, docs, counts)."""
        nq = int(scores.shape[0])
        assert 0 < nq <= self.nq and scores.shape[1] == self.k
        par = self.step & 1
        self.step += 1
        slots = [self._slot_ptrs(self.ptrs[r], par, self.rank) for r in range(self.world)]
        ops.exchange_publish(scores, docs, counts, [p[0] for p in slots], [p[1] for p in slots], [p[2] for p in slots])
        s0, d0, c0 = self._slot_ptrs(self.ptrs[self.rank], par, 0)
        final = self.final if nq == self.nq else tuple(t[:nq] for t in self.final)
        ops.topk_merge_exchange(s0, d0, c0, self.world, nq, self.k, self.slot, self.slot, self.k, final, self.device,
                                self.rank, self.ptrs[self.rank] + 4 * self.flag_off,
                                [self.ptrs[r] + 4 * self.flag_off for r in self.peers], self.peers, self.step)
        return final

    def phase_ms(self, last=None):
        """(search ms, exchange-merge ms incl. the wait for the slowest peer) averaged over the last ``last`` profiled steps."""
        torch.cuda.synchronize(self.device)
        ev = self._ev[-last:] if last else self._ev
        self._ev = []
        if not ev:
            return 0.0, 0.0
        return (sum(e[0].elapsed_time(e[1]) for e in ev) / len(ev), sum(e[1].elapsed_time(e[2]) for e in ev) / len(ev))


def make_peer_exchange(nq, k, device, world, rank, kind="auto", group=None):
    """``PeerExchange`` if every rank of the group can set it up, else None (``kind='auto'``) or an error (``'p2p'``);
    ``kind='nccl'`` returns None at once.  Collective: every rank must call it with the same arguments."""
    if world == 1 or kind == "nccl":
        return None
    px, err = None, None
    try:
        px = PeerExchange(nq, k, device, world, rank, group=group)
    except Exception as e:                                          # no symmetric memory / no peer access
        err = e
    ok = torch.tensor([1 if px is not None else 0], device=torch.device(device))
    dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=group)
    if int(ok.item()) == 0:
        if kind == "p2p":
            raise RuntimeError(f"fused peer exchange unavailable on rank {rank}: {err!r}") from err
        return None
    return px
