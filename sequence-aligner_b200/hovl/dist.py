"""hovl.dist — the hash-overlap path over several GPUs.

One process per GPU (torch.distributed, NCCL over NVLink/NVSwitch).  The path shards with two
real exchange steps, so there are exactly two data-path collectives plus the replication of the
2-bit packed reads and the final gather of records:

  reads by id  ->  splitter on the local reads              Context.shard_encode
               ->  ALL-TO-ALL #1: k-mer tuples to the rank that owns mix(code)
               ->  KmerTable slice + partial PairData        Context.shard_pairs
               ->  ALL-TO-ALL #2: (pair key, count) to the rank that owns the pair's fst read
               ->  sum, count filter, dovetail DP            Context.shard_align
               ->  records gathered on rank 0 (rank order == canonical (a, b) order)

The exchange layer below only needs torch tensors and a process group, so it runs unchanged
on gloo/CPU; tests/test_dist_cpu.py drives it with the oracle standing in for the kernels.
"""
import os
import sys
import time

import numpy as np
import torch
import torch.distributed as dist

STAT_SUM = ("n_reads", "n_bases", "n_kmers", "n_buckets", "n_pair_instances", "n_distinct_pairs", "n_dispatched",
            "n_leads", "n_duds", "n_align_valid", "n_ovl_valid", "dp_cells", "n_records", "gpu_launches",
            "n_list_entries", "dp_cells_skipped", "dp_cells_phase1")
REBALANCE_OVER = 1.08      # pairs move when the heaviest rank's DP cost exceeds this multiple of the mean

STAT_MAX = ("max_bucket", "ns_h2d", "ns_pack", "ns_encode", "ns_kmer_sort", "ns_group", "ns_expand", "ns_pair_sort",
            "ns_pair_reduce", "ns_order", "ns_dp", "ns_d2h", "ns_total")


def read_slice(n_reads, rank, world):
    """Contiguous slice of read indices (0-based, half open) a rank encodes."""
    return n_reads * rank // world, n_reads * (rank + 1) // world


# ------------------------------------------------------------------ exchange layer (backend-agnostic)
def exchange_counts(send_counts, device, group=None):
    """send_counts[p] = items this rank has for rank p  ->  recv_counts[p] = items rank p has for us."""
    t = torch.tensor(send_counts, dtype=torch.int64, device=device)
    r = torch.empty_like(t)
    dist.all_to_all_single(r, t, group=group)
    return [int(x) for x in r.tolist()]


def all_to_all_v(send, send_counts, recv_counts, group=None):
    """Variable all-to-all of a 1-D tensor laid out as one contiguous segment per destination."""
    recv = torch.empty(int(sum(recv_counts)), dtype=send.dtype, device=send.device)
    dist.all_to_all_single(recv, send[:int(sum(send_counts))], output_split_sizes=list(recv_counts),
                           input_split_sizes=list(send_counts), group=group)
    return recv


def all_gather_v(t, group=None):
    """Concatenation of every rank's 1-D tensor in rank order (sizes may differ) + the sizes."""
    world = dist.get_world_size(group)
    n = torch.tensor([t.numel()], dtype=torch.int64, device=t.device)
    sizes = [torch.empty_like(n) for _ in range(world)]
    dist.all_gather(sizes, n, group=group)
    sizes = [int(s.item()) for s in sizes]
    m = max(sizes) if sizes else 0
    pad = torch.zeros(m, dtype=t.dtype, device=t.device)
    pad[:t.numel()] = t
    parts = [torch.empty_like(pad) for _ in range(world)]
    dist.all_gather(parts, pad, group=group)
    return torch.cat([p[:s] for p, s in zip(parts, sizes)]), sizes


def all_gather_multi(ts, group=None):
    """Several 1-D tensors of one dtype per rank -> for each of them the concatenation over the ranks (rank order).
    Two collectives and one host sync for the whole set (all_gather_v costs two collectives and `world` syncs each)."""
    world, dev, k = dist.get_world_size(group), ts[0].device, len(ts)
    n = torch.tensor([int(t.numel()) for t in ts], dtype=torch.int64, device=dev)
    alln = torch.empty(world * k, dtype=torch.int64, device=dev)
    dist.all_gather_into_tensor(alln, n, group=group)
    sizes = alln.view(world, k).tolist()
    m = max(1, max(sum(r) for r in sizes))
    pad = torch.zeros(m, dtype=ts[0].dtype, device=dev)
    o = 0
    for t in ts:
        pad[o:o + int(t.numel())] = t
        o += int(t.numel())
    out = torch.empty(world * m, dtype=ts[0].dtype, device=dev)
    dist.all_gather_into_tensor(out, pad, group=group)
    out = out.view(world, m)
    res = []
    for j in range(k):
        res.append(torch.cat([out[r, sum(sizes[r][:j]):sum(sizes[r][:j + 1])] for r in range(world)]))
    return res


def all_gather_sized(ts, sizes, group=None):
    """all_gather_multi when every rank already knows every rank's sizes (sizes[r][j] = elements of tensor j on rank r):
    one collective, no host synchronisation."""
    world, dev, k = dist.get_world_size(group), ts[0].device, len(ts)
    m = max(1, max(sum(r) for r in sizes))
    pad = torch.zeros(m, dtype=ts[0].dtype, device=dev)
    o = 0
    for t in ts:
        pad[o:o + int(t.numel())] = t
        o += int(t.numel())
    out = torch.empty(world * m, dtype=ts[0].dtype, device=dev)
    dist.all_gather_into_tensor(out, pad, group=group)
    out = out.view(world, m)
    res = []
    for j in range(k):
        res.append(torch.cat([out[r, sum(sizes[r][:j]):sum(sizes[r][:j + 1])] for r in range(world)]))
    return res


def gather_v(t, dst=0, group=None):
    """Rank dst gets the concatenation in rank order; the others get None.  One variable
    all-to-all in which only dst receives, so nothing is padded or replicated."""
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    send = [0] * world
    send[dst] = int(t.numel())
    recv = exchange_counts(send, t.device, group)
    out = all_to_all_v(t, send, recv, group)
    return out if rank == dst else None


def reduce_stats(stats, device, group=None):
    """Whole-job statistics: counters summed over ranks, stage times as the max over ranks."""
    s = torch.tensor([stats[k] for k in STAT_SUM], dtype=torch.int64, device=device)
    m = torch.tensor([stats[k] for k in STAT_MAX], dtype=torch.int64, device=device)
    dist.all_reduce(s, op=dist.ReduceOp.SUM, group=group)
    dist.all_reduce(m, op=dist.ReduceOp.MAX, group=group)
    out = dict(stats)
    out.update({k: int(v) for k, v in zip(STAT_SUM, s.tolist())})
    out.update({k: int(v) for k, v in zip(STAT_MAX, m.tolist())})
    return out


class PeerExchange:
    """Receive buffers of all-to-all #1 that the other ranks of the node store into directly: every rank owns a
    (codes, vals) pair of device buffers allocated through the C ABI, hands their CUDA IPC handles to the others and
    maps theirs.  Capacities are tracked for ALL ranks on every rank (they follow from the count matrix everybody
    sees), so growing a buffer is a collective decision without an extra message."""

    def __init__(self, ctx, group=None):
        self.ctx, self.group = ctx, group
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
        self.caps = [0] * self.world
        self.own = None                                  # (codes address, vals address) of this rank's buffers
        self.codes, self.vals = [0] * self.world, [0] * self.world   # every rank's buffers as mapped HERE

    def ensure(self, totals, dev):
        grow = [t > c for t, c in zip(totals, self.caps)]
        if not any(grow):
            return
        for r in range(self.world):                      # drop the mappings of buffers that are about to go
            if grow[r] and r != self.rank and self.codes[r]:
                self.ctx.peer_close(self.codes[r])
                self.ctx.peer_close(self.vals[r])
                self.codes[r] = self.vals[r] = 0
        dist.barrier(group=self.group)
        handles = torch.zeros(128, dtype=torch.uint8)
        error = None                                     # a local failure must not skip the collectives below
        if grow[self.rank]:
            try:
                if self.own:
                    self.ctx.peer_free(self.own[0])
                    self.ctx.peer_free(self.own[1])
                    self.own = None
                cap = int(totals[self.rank] * 1.25) + 4096
                pc, hc = self.ctx.peer_alloc(cap * 4)
                pv, hv = self.ctx.peer_alloc(cap * 8)
                self.own = (pc, pv)
                self.codes[self.rank], self.vals[self.rank] = pc, pv
                handles = torch.tensor(list(hc + hv), dtype=torch.uint8)
            except Exception as e:
                error = e
        for r in range(self.world):
            if grow[r]:
                self.caps[r] = int(totals[r] * 1.25) + 4096
        allh = torch.empty(self.world * 128, dtype=torch.uint8, device=dev)
        dist.all_gather_into_tensor(allh, handles.to(dev), group=self.group)
        allh = allh.cpu().numpy().reshape(self.world, 128)
        for r in range(self.world):
            if grow[r] and r != self.rank:
                try:
                    if not allh[r].any():
                        raise RuntimeError("rank %d could not export its buffers" % r)
                    self.codes[r] = self.ctx.peer_open(bytes(allh[r, :64]))
                    self.vals[r] = self.ctx.peer_open(bytes(allh[r, 64:]))
                except Exception as e:
                    error = error or e
        if error is not None:
            raise error

    def close(self):
        """Collective: mappings go first, then (after a barrier) the buffers themselves."""
        for r in range(self.world):
            if r != self.rank and self.codes[r]:
                self.ctx.peer_close(self.codes[r])
                self.ctx.peer_close(self.vals[r])
        dist.barrier(group=self.group)
        if self.own:
            self.ctx.peer_free(self.own[0])
            self.ctx.peer_free(self.own[1])
        self.own, self.codes, self.vals, self.caps = None, [0] * self.world, [0] * self.world, [0] * self.world


_REGISTERED = {}          # host ranges this process has page-locked with cudaHostRegister: address -> bytes


def _host_unregister(ptr):
    if _REGISTERED.pop(ptr, None) is not None:
        try:
            torch.cuda.cudart().cudaHostUnregister(ptr)
        except Exception:
            pass


class SharedLanding:
    """Where the records of a one-node job land: a host buffer shared by the ranks (a file in /dev/shm mapped by
    every rank and page-locked there with cudaHostRegister).  Every rank copies ITS slice device -> host over its own
    PCIe link, at the offset its rank order gives it, so rank 0 holds the whole job's records without first pulling
    them through one GPU and one link (143 MB on K3: 2.5 ms of D2H on rank 0, 0.3 ms when eight links share it)."""

    def __init__(self, group=None):
        self.group, self.cap, self.t, self.mm = group, 0, None, None
        self.rank = dist.get_rank(group)
        self.serial = 0
        self._same_host = None

    def same_host(self, dev):
        """Collective, once: do all ranks of the group run on one machine (same hostname and boot id)?"""
        if self._same_host is None:
            boot = ""
            try:
                with open("/proc/sys/kernel/random/boot_id") as f:
                    boot = f.read().strip()
            except OSError:
                pass
            me = (os.uname().nodename, boot, os.path.isdir("/dev/shm"))
            world = dist.get_world_size(self.group)
            everyone = [None] * world
            dist.all_gather_object(everyone, me, group=self.group)
            self._same_host = all(e == everyone[0] for e in everyone) and bool(me[2])
        return self._same_host

    def ensure(self, nbytes, dev):
        """Collective: every rank calls it with the same nbytes.  Returns True when the buffer is mapped on EVERY rank;
        on any failure (several nodes, /dev/shm too small, mmap refused) every rank gets False together and the caller
        falls back to the gather through rank 0's GPU."""
        if not self.same_host(dev):
            return False
        if nbytes <= self.cap:
            return True
        self.close()
        import mmap
        cap = ((int(nbytes * 1.25) + (4 << 20)) + (2 << 20) - 1) // (2 << 20) * (2 << 20)
        name, ok = [None], 1
        if self.rank == 0:
            self.serial += 1
            path = "/dev/shm/hovl_records_%d_%d" % (os.getpid(), self.serial)
            try:
                fd = os.open(path, os.O_CREAT | os.O_EXCL | os.O_RDWR, 0o600)
                try:
                    os.posix_fallocate(fd, 0, cap)       # reserve the pages now: ftruncate alone reserves nothing on tmpfs,
                finally:                                 # and a later copy into an over-committed mapping dies with SIGBUS
                    os.close(fd)
                name[0] = path
            except OSError as e:
                print("[hovl.dist] no shared landing buffer (%s); records go through rank 0's GPU" % e, file=sys.stderr, flush=True)
                try:
                    os.unlink(path)
                except OSError:
                    pass
        src = dist.get_global_rank(self.group, 0) if self.group is not None else 0
        dist.broadcast_object_list(name, src=src, group=self.group)
        if name[0] is None:
            ok = 0
        else:
            try:
                fd = os.open(name[0], os.O_RDWR)
                try:
                    self.mm = mmap.mmap(fd, cap)
                finally:
                    os.close(fd)
                self.t = torch.frombuffer(self.mm, dtype=torch.uint8)
            except (OSError, ValueError) as e:
                print("[hovl.dist] rank %d cannot map the shared landing buffer (%s)" % (self.rank, e), file=sys.stderr, flush=True)
                ok, self.mm, self.t = 0, None, None
        flag = torch.tensor([ok], dtype=torch.int32, device=dev)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=self.group)
        all_ok = int(flag.item())                # the host waits here (an NCCL collective only queues work): every rank has
        if self.rank == 0 and name[0] is not None:   # tried to map the file by now, so the name can go
            try:
                os.unlink(name[0])
            except OSError:
                pass
        if all_ok == 0:
            self.close()
            return False
        ptr = self.t.data_ptr()
        for p, c in list(_REGISTERED.items()):   # a mapping that is gone must not leave its registration behind
            if p < ptr + cap and ptr < p + c:
                _host_unregister(p)
        rc = torch.cuda.cudart().cudaHostRegister(ptr, cap, 0)
        code = int(getattr(rc, "value", rc))
        self.registered = code == 0
        if self.registered:
            _REGISTERED[ptr] = cap
        else:                                    # still correct (the pages are reserved): copies go through the driver's staging buffers
            print("[hovl.dist] cudaHostRegister of the shared record buffer failed (error %d); using it unpinned" % code,
                  file=sys.stderr, flush=True)
        self.cap = cap
        return True

    def close(self):
        if self.t is not None:
            if getattr(self, "registered", False):
                _host_unregister(self.t.data_ptr())
                self.registered = False
            self.t = None
        if self.mm is not None:
            try:
                self.mm.close()
            except (BufferError, ValueError):
                pass                             # a caller still holds a view of the last result
            self.mm = None
        self.cap = 0

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


# ------------------------------------------------------------------ the sharded overlapper
class ShardedResult:
    """records (rank 0 only) is a VIEW of the reused landing buffer: it is valid until the next run_device / run_host /
    close of the same ShardedOverlapper.  Pass copy=True to run_device / run_host (or call owned()) for an array of
    its own that outlives them."""

    def owned(self):
        """Detaches records from the landing buffer (one host copy); returns self."""
        if self.records is not None:
            self.records = np.array(self.records)
        return self

    def __init__(self, records, stats, local_stats, ktimes):
        self.records, self.stats, self.local_stats, self.ktimes = records, stats, local_stats, ktimes


class GpuStages:
    """The three compute stages of a rank on its GPU: thin wrappers over the C ABI
    (hovl_shard_encode / hovl_shard_pairs + hovl_shard_fetch_pairs / hovl_shard_align) that
    take and return CUDA tensors."""

    def __init__(self, ctx, settings, stream=None):
        self.ctx, self.s, self.stream = ctx, settings, stream

    def _sync(self, dev):
        # the C calls launch on the context's stream: everything torch has queued for them (on the
        # current stream and on ours) must have landed first
        torch.cuda.current_stream(dev).synchronize()
        if self.stream is not None:
            self.stream.synchronize()

    def encode(self, d_bases, d_offsets, lo, n_total, parts, keep=False):
        """keep=True: the tuples stay in the context for scatter() (peer-memory exchange); codes / vals are None."""
        dev, n_local, n_bases = d_bases.device, int(d_offsets.numel()) - 1, int(d_bases.numel())
        tuple_cap, packed_cap = max(n_bases, 1), n_bases // 16 + n_local + 2
        codes = None if keep else torch.empty(tuple_cap, dtype=torch.int32, device=dev)
        vals = None if keep else torch.empty(tuple_cap, dtype=torch.int64, device=dev)
        packed = torch.empty(packed_cap, dtype=torch.int32, device=dev)
        lens = torch.empty(max(n_local, 1), dtype=torch.int32, device=dev)
        bad = torch.empty(max(n_local, 1), dtype=torch.int32, device=dev)
        self._sync(dev)
        counts, n_words = self.ctx.shard_encode(self.s, d_bases.data_ptr(), d_offsets.data_ptr(), n_local, n_bases, lo,
                                                n_total, parts, 0 if keep else codes.data_ptr(), 0 if keep else vals.data_ptr(), tuple_cap,
                                                packed.data_ptr(), packed_cap, lens.data_ptr(), bad.data_ptr())
        return codes, vals, counts, packed[:n_words], lens[:n_local], bad[:n_local]

    def scatter(self, parts, peer_codes, peer_vals, peer_base):
        self.ctx.shard_scatter(parts, peer_codes, peer_vals, peer_base)

    def pairs(self, r_codes, r_vals, n_total, parts, n_tuples=None, dev=None):
        """r_codes / r_vals: tensors, or raw device addresses together with n_tuples and dev."""
        if n_tuples is None:
            dev, n_tuples, r_codes, r_vals = r_codes.device, int(r_codes.numel()), r_codes.data_ptr(), r_vals.data_ptr()
        self._sync(dev)
        counts, n_partial = self.ctx.shard_pairs(self.s, r_codes, r_vals, n_tuples, n_total, parts)
        keys = torch.empty(max(n_partial, 1), dtype=torch.int64, device=dev)
        cnts = torch.empty(max(n_partial, 1), dtype=torch.int32, device=dev)
        self._sync(dev)
        self.ctx.shard_fetch_pairs(keys.data_ptr(), cnts.data_ptr(), n_partial)
        return keys[:n_partial], cnts[:n_partial], counts

    def align(self, r_keys, r_cnts, all_packed, all_len, all_bad, n_total):
        import hovl
        dev, cap = r_keys.device, max(int(r_keys.numel()), 1)
        recs = torch.empty(cap * hovl.RECORD_DTYPE.itemsize, dtype=torch.uint8, device=dev)
        self._sync(dev)
        res = self.ctx.shard_align(self.s, r_keys.data_ptr(), r_cnts.data_ptr(), int(r_keys.numel()),
                                   all_packed.data_ptr(), int(all_packed.numel()), all_len.data_ptr(),
                                   all_bad.data_ptr(), n_total, recs.data_ptr(), cap)
        return recs[:res.stats["n_records"] * hovl.RECORD_DTYPE.itemsize], res.stats   # records stay on the device

    # ---- alignment re-balance by DP cells
    def select(self, r_keys, r_cnts, all_len, n_total):
        """Sum + count filter; the dispatched pairs stay in the context.  Returns (pairs, their DP cost in cells)."""
        self._sync(r_keys.device)
        return self.ctx.shard_select(self.s, r_keys.data_ptr(), r_cnts.data_ptr(), int(r_keys.numel()), all_len.data_ptr(), n_total)

    def split(self, cost_before, cost_total, parts):
        return self.ctx.shard_split(cost_before, cost_total, parts)

    def fetch_selected(self, n_pairs, dev):
        abc = torch.empty(max(3 * n_pairs, 1), dtype=torch.int32, device=dev)
        self._sync(dev)
        self.ctx.shard_fetch_selected(abc.data_ptr(), n_pairs)
        return abc[:3 * n_pairs]

    def align_pairs(self, abc, all_packed, all_len, all_bad, n_total):
        import hovl
        dev, n = abc.device, int(abc.numel()) // 3
        recs = torch.empty(max(n, 1) * hovl.RECORD_DTYPE.itemsize, dtype=torch.uint8, device=dev)
        self._sync(dev)
        res = self.ctx.shard_align_pairs(self.s, abc.data_ptr(), n, all_packed.data_ptr(), int(all_packed.numel()),
                                         all_len.data_ptr(), all_bad.data_ptr(), n_total, recs.data_ptr(), max(n, 1))
        return recs[:res.stats["n_records"] * hovl.RECORD_DTYPE.itemsize], res.stats

    # ---- replicated-table variant
    def sort(self, r_codes, r_vals, n_tuples=None, dev=None):
        """Stable sort of the received tuples by code, in place (tensors, or raw addresses with n_tuples and dev)."""
        if n_tuples is None:
            dev, n_tuples, r_codes, r_vals = r_codes.device, int(r_codes.numel()), r_codes.data_ptr(), r_vals.data_ptr()
        self._sync(dev)
        self.ctx.shard_sort(self.s, r_codes, r_vals, n_tuples)

    def gather_peer(self, parts, peer_codes, peer_vals, counts, dev):
        n_all = int(sum(counts))
        codes = torch.empty(max(n_all, 1), dtype=torch.int32, device=dev)
        vals = torch.empty(max(n_all, 1), dtype=torch.int64, device=dev)
        self._sync(dev)
        self.ctx.shard_gather_peer(parts, peer_codes, peer_vals, counts, codes.data_ptr(), vals.data_ptr())
        return codes[:n_all], vals[:n_all]

    def pairs_local(self, all_codes, all_vals, all_packed, all_len, all_bad, n_total, lo, n_local):
        dev = all_codes.device
        self._sync(dev)
        res = self.ctx.shard_pairs_local(self.s, all_codes.data_ptr(), all_vals.data_ptr(), int(all_codes.numel()),
                                         all_packed.data_ptr(), int(all_packed.numel()), all_len.data_ptr(),
                                         all_bad.data_ptr(), n_total, lo, n_local)
        return None, res.stats, self.ctx.shard_fetch_records           # the records wait on the device for fetch(dst, cap)

    def kernel_times(self):
        return self.ctx.kernel_times()


class ShardedOverlapper:
    """generateKmerTable + genAlignment over `world` ranks.  Every rank calls run_device with its
    slice of the reads; rank 0's result holds all records in canonical order."""

    def __init__(self, ctx, settings, stream=None, group=None, stages=None, mode="partials", rebalance=True):
        """mode "partials": the k-mer table stays sharded by a mixed hash of the code and PARTIAL pair counts travel to
        the rank that owns the pair's fst read (all-to-all #2).  mode "replicate": ranks own code RANGES, the sorted
        slices are gathered on every rank and each rank counts the pairs of its own reads against the whole table
        (include/hovl.h, "replicated-table variant") — for tables dominated by chance matches (K4)."""
        if settings.id_mode != 1 or settings.order != 1:
            raise ValueError("the sharded path needs id_mode=1, order=1 (canonical order, wide pair keys)")
        if mode not in ("partials", "replicate"):
            raise ValueError("mode must be 'partials' or 'replicate'")
        self.mode = mode
        self.rebalance = bool(rebalance) and os.environ.get("HOVL_DIST_REBALANCE", "1") != "0"   # mode "partials" only
        self.last_rebalance = None
        self.stages = stages if stages is not None else GpuStages(ctx, settings, stream)
        if stages is None:
            ctx.set_option("shard_by_range", 1 if mode == "replicate" else 0)
        elif hasattr(stages, "by_range"):
            stages.by_range = mode == "replicate"
        self.settings, self.stream, self.group = settings, stream, group
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
        self._pinned = None      # rank 0: page-locked landing buffer for the gathered records (reused)
        # one node, CUDA: records land in a host buffer shared by the ranks (HOVL_DIST_SHARED_LANDING=0 turns it off)
        self._landing = None
        self._use_landing = os.environ.get("HOVL_DIST_SHARED_LANDING", "1") != "0"
        # one node, CUDA: all-to-all #1 is the partition kernel storing into the peers' buffers (HOVL_DIST_PEER_SCATTER=0: NCCL)
        self._peer = None
        self._use_peer = os.environ.get("HOVL_DIST_PEER_SCATTER", "1") != "0" and stages is None and self.world <= 32

    def close(self):
        """Collective: releases the peer-mapped exchange buffers and the shared landing buffer."""
        if self._peer is not None:
            self._peer.close()
            self._peer = None
        if self._landing is not None:
            self._landing.close()
            self._landing = None

    def run_host(self, h_bases, lo, hi, n_total, read_len, copy=False):
        dev = torch.device("cuda", torch.cuda.current_device())
        with torch.cuda.stream(self.stream or torch.cuda.current_stream(dev)):
            d = h_bases.to(dev, non_blocking=True)
            return self.run_device(d, lo, hi, n_total, read_len, copy=copy)

    def run_device(self, d_bases, lo, hi, n_total, read_len=None, d_offsets=None, copy=False):
        """d_bases: uint8 tensor with this rank's reads [lo, hi) concatenated; fixed read_len, or
        explicit local offsets (int64 tensor, hi-lo+1 entries).  The result's records are a view of a buffer the next
        run reuses unless copy=True (see ShardedResult)."""
        dev = d_bases.device

        def go():
            offs = d_offsets
            if offs is None:
                offs = torch.arange(hi - lo + 1, dtype=torch.int64, device=dev) * read_len
            res = self._run(d_bases, offs, lo, n_total)
            return res.owned() if copy else res
        if dev.type == "cuda":
            with torch.cuda.stream(self.stream or torch.cuda.current_stream(dev)):
                return go()
        return go()

    def _run(self, d_bases, d_offsets, lo, n_total):
        import hovl
        S, G, grp, dev = self.stages, self.world, self.group, d_bases.device
        trace = [] if os.environ.get("HOVL_DIST_TRACE") else None
        replicate = self.mode == "replicate"
        n_local = int(d_offsets.numel()) - 1

        def mark(name):
            if trace is not None:
                if dev.type == "cuda":
                    torch.cuda.synchronize(dev)
                trace.append((name, time.perf_counter()))
        mark("start")
        peer = dev.type == "cuda" and self._use_peer
        codes, vals, send_counts, packed, lens, bad = S.encode(d_bases, d_offsets, lo, n_total, G, **({"keep": True} if peer else {}))
        mark("encode")
        # ---- all-to-all #1: k-mer tuples to the owner of the code.  One all-gather carries the G x G count matrix AND the
        # sizes of the read arrays every rank is about to replicate, so their all-gather needs no size exchange of its own.
        row = torch.tensor(list(send_counts) + [int(packed.numel()), int(lens.numel())], dtype=torch.int64, device=dev)
        mat = torch.empty(G * (G + 2), dtype=torch.int64, device=dev)
        dist.all_gather_into_tensor(mat, row, group=grp)
        mat = mat.view(G, G + 2).tolist()                 # mat[r][p]: tuples rank r has for rank p; then words, reads of rank r
        totals = [sum(mat[r][p] for r in range(G)) for p in range(G)]
        read_sizes = [[mat[r][G], mat[r][G + 1], mat[r][G + 1]] for r in range(G)]
        if peer:
            base = [sum(mat[r][p] for r in range(self.rank)) for p in range(G)]
            if self._peer is None:
                self._peer = PeerExchange(S.ctx, grp)
            grew = any(t > c for t, c in zip(totals, self._peer.caps))
            ok = 1
            try:
                self._peer.ensure(totals, dev)
            except Exception as e:                        # no IPC / no peer access here: every rank falls back together
                ok = 0
                print("[hovl.dist] peer-memory exchange unavailable on rank %d (%s); using NCCL" % (self.rank, e),
                      file=sys.stderr, flush=True)
            if grew:                                      # mapping happens only then, so only then can it fail
                flag = torch.tensor([ok], dtype=torch.int32, device=dev)
                dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=grp)
                if int(flag.item()) == 0:
                    self._use_peer, peer = False, False
                    codes, vals, send_counts, packed, lens, bad = S.encode(d_bases, d_offsets, lo, n_total, G)
        r_codes = r_vals = None
        if peer:
            # the count matrix told every rank where its segment starts in every receive buffer; the partition kernel
            # now stores each tuple straight into its owner's buffer (peer memory over NVLink)
            S.scatter(G, self._peer.codes, self._peer.vals, base)
            mark("all_to_all_1")
            if not replicate:
                # the all-gather of the reads is also the barrier after which every rank's stores have landed
                all_packed, all_len, all_bad = all_gather_sized([packed, lens, bad], read_sizes, grp)
                mark("all_gather_reads")
        else:
            recv_counts = [mat[r][self.rank] for r in range(G)]
            r_codes = all_to_all_v(codes, send_counts, recv_counts, grp)
            r_vals = all_to_all_v(vals, send_counts, recv_counts, grp)
            del codes, vals
            mark("all_to_all_1")
            if not replicate:
                # ---- replicate the packed reads (2 bit/base), their lengths and bad-base flags
                all_packed, all_len, all_bad = all_gather_sized([packed, lens, bad], read_sizes, grp)
                mark("all_gather_reads")
        fetch = None
        if replicate:
            # ---- replicated-table variant: sort the code range where it landed, gather every rank's sorted slice (in rank
            # order they are the sorted table), then count + filter + align the pairs whose fst is one of OUR reads
            if peer:
                dist.barrier(group=grp)                   # every rank's stores into our buffer have landed
                S.sort(self._peer.own[0], self._peer.own[1], totals[self.rank], dev=dev)
            else:
                S.sort(r_codes, r_vals)
            mark("sort")
            # the all-gather of the reads doubles as the barrier after which every slice is sorted
            all_packed, all_len, all_bad = all_gather_sized([packed, lens, bad], read_sizes, grp)
            mark("all_gather_reads")
            if peer:
                all_codes, all_vals = S.gather_peer(G, self._peer.codes, self._peer.vals, totals, dev)
            else:
                all_codes, all_vals = all_gather_sized([r_codes], [[t] for t in totals], grp)[0], \
                    all_gather_sized([r_vals], [[t] for t in totals], grp)[0]
                del r_codes, r_vals
            mark("table_gather")
            records, stats, fetch = S.pairs_local(all_codes, all_vals, all_packed, all_len, all_bad, n_total, lo, n_local)
            del all_codes, all_vals
            ktimes = S.kernel_times()
            mark("pairs_local")
        else:
            if peer:
                keys, cnts, pair_counts = S.pairs(self._peer.own[0], self._peer.own[1], n_total, G, n_tuples=totals[self.rank], dev=dev)
            else:
                keys, cnts, pair_counts = S.pairs(r_codes, r_vals, n_total, G)
                del r_codes, r_vals
            mark("pairs")
            # ---- all-to-all #2: partial (pair, count) to the owner of fst
            recv_pairs = exchange_counts(pair_counts, dev, grp)
            r_keys = all_to_all_v(keys, pair_counts, recv_pairs, grp)
            r_cnts = all_to_all_v(cnts, pair_counts, recv_pairs, grp)
            del keys, cnts
            mark("all_to_all_2")
            if self.rebalance and hasattr(S, "select"):
                # ---- alignment re-balance by DP cells (north_star): every rank tells what its dispatched pairs cost; when the
                # heaviest rank carries more than REBALANCE_OVER x the mean, pairs move so that every rank aligns an equal share
                n_pairs, cost = S.select(r_keys, r_cnts, all_len, n_total)
                del r_keys, r_cnts
                mine = torch.tensor([cost], dtype=torch.int64, device=dev)
                costs = torch.empty(G, dtype=torch.int64, device=dev)
                dist.all_gather_into_tensor(costs, mine, group=grp)
                costs = costs.tolist()
                total = sum(costs)
                moved = total > 0 and max(costs) * G > REBALANCE_OVER * total
                seg = S.split(sum(costs[:self.rank]), total, G) if moved else None     # before the fetch: it reads the cost prefix
                abc = S.fetch_selected(n_pairs, dev)
                if moved:
                    assert sum(seg) == n_pairs, (seg, n_pairs)
                    got = exchange_counts(seg, dev, grp)
                    abc = all_to_all_v(abc, [3 * x for x in seg], [3 * x for x in got], grp)
                self.last_rebalance = {"cost": costs, "moved": bool(moved)}
                mark("rebalance")
                records, stats = S.align_pairs(abc, all_packed, all_len, all_bad, n_total)
            else:
                records, stats = S.align(r_keys, r_cnts, all_packed, all_len, all_bad, n_total)
            ktimes = S.kernel_times()
            mark("align")
        # ---- records to rank 0 (rank order == canonical order: fst ranges ascend with the rank)
        rec_size = hovl.RECORD_DTYPE.itemsize
        if fetch is not None:
            mine, my_bytes = None, int(stats["n_records"]) * rec_size
        else:
            mine = records if torch.is_tensor(records) else torch.from_numpy(np.ascontiguousarray(records).view(np.uint8).copy())
            my_bytes = int(mine.numel())
        gathered, out, landed = None, None, False
        if dev.type == "cuda" and self._use_landing:
            # collective 1: every rank's record bytes -> offsets in the node's shared page-locked buffer
            sz = torch.tensor([my_bytes], dtype=torch.int64, device=dev)
            szs = torch.empty(G, dtype=torch.int64, device=dev)
            dist.all_gather_into_tensor(szs, sz, group=grp)
            sizes = szs.tolist()
            mark("records_sizes")
            total, off = sum(sizes), sum(sizes[:self.rank])
            if self._landing is None:
                self._landing = SharedLanding(grp)
            ok = self._landing.ensure(max(total, 1), dev)          # collective; False on every rank if any rank failed
            if not ok:
                self._use_landing = False
            else:
                if my_bytes:
                    if fetch is not None:
                        fetch(self._landing.t.data_ptr() + off, my_bytes // rec_size)
                    else:
                        self._landing.t[off:off + my_bytes].copy_(mine, non_blocking=True)
                torch.cuda.current_stream(dev).synchronize()      # my slice has landed ...
                mark("records_copy")
                # collective 2: the statistics rows; it is queued behind the copy on every rank, so once it returns
                # every rank's slice has landed as well
                row = torch.tensor([stats[k] for k in STAT_SUM] + [stats[k] for k in STAT_MAX], dtype=torch.int64, device=dev)
                rows = torch.empty(G * row.numel(), dtype=torch.int64, device=dev)
                dist.all_gather_into_tensor(rows, row, group=grp)
                rows = rows.view(G, row.numel()).tolist()
                whole = dict(stats)
                whole.update({k: sum(r[i] for r in rows) for i, k in enumerate(STAT_SUM)})
                whole.update({k: max(r[len(STAT_SUM) + i] for r in rows) for i, k in enumerate(STAT_MAX)})
                landed = True
        if not landed:
            if fetch is not None:                       # records still wait in the context: bring them into a tensor first
                mine = torch.empty(max(my_bytes, 1), dtype=torch.uint8, device=dev)[:my_bytes]
                if my_bytes:
                    fetch(mine.data_ptr(), my_bytes // rec_size)
            gathered = gather_v(mine.to(dev), 0, grp)
            whole = reduce_stats(stats, dev, grp)
        if landed and self.rank == 0:
            out = self._landing.t[:total].numpy().view(hovl.RECORD_DTYPE)
        if gathered is not None:
            if gathered.is_cuda:                     # one device -> pinned host copy of the whole job's records
                if self._pinned is None or self._pinned.numel() < gathered.numel():
                    self._pinned = torch.empty(int(gathered.numel() * 1.25) + 4096, dtype=torch.uint8, pin_memory=True)
                host = self._pinned[:gathered.numel()]
                host.copy_(gathered, non_blocking=False)
                out = host.numpy().view(hovl.RECORD_DTYPE)
            else:
                out = gathered.numpy().view(hovl.RECORD_DTYPE)
        mark("gather_records")
        if trace is not None and self.rank == 0:
            print("[hovl.dist] " + "  ".join("%s %.1f ms" % (b[0], 1e3 * (b[1] - a[1])) for a, b in zip(trace, trace[1:])),
                  file=sys.stderr, flush=True)
        return ShardedResult(out, whole, stats, ktimes)
