#!/usr/bin/env python
"""bench.py — reference Gbp k-mer-screened per second on B200 (BASELINE.json's metric).

A step = one pass of the whole `eref` hot path over one synthetic sample: reset the 2^32-slot table,
frame+pack+count both FASTQ files, frame+pack the FASTA database, scan it (candidate filter, probes,
500-position window rule, interval merge) down to one {ordinal, frag, el, ref_len} record per
reference and, at N>1, all-gather the records.

  value   inputs already resident in HBM (eref_*_device entry points), CUDA events, max over ranks
  e2e     the same through the host-buffer C ABI call (pinned host -> H2D -> ... -> D2H of the records)

The ONE JSON line carries, besides the contract's keys:
  headline   BASELINE configs[1] per GPU (weak scaling: every rank its own 10,000-genome ~500 Mbp shard, the
             50 Mbp query replicated) — `value`, `e2e`, `phases`, `roofline`
  strong     BASELINE configs[2] (and [3] with --strong c3,c4): ONE 30,000-genome ~1.5 Gbp database sharded over the
             N GPUs, 50 Mbp (500 Mbp) of reads — the strong-scaling curve north_star's target is written for
  c5         BASELINE configs[4]: short-read-heavy query (20-150 bp reads) against the headline database
  scan_unpruned   the scan with the candidate filter off (all 3 probes at every position), the only figure
             comparable with SURVEY 8(d)'s 97 B per reference base
  parity     the records of the workloads that were timed, compared with the CPU oracle in this very run
  cpu_baseline    (N=1) the unmodified reference on the host cores, bounded sample, and the whole-process comparison
At N>1 the query-table build is sharded over the GPUs (each counts 1/N of the table slices and stores them into
every peer's table over NVLink, include/eref_gpu.h); the DB is sharded by record; the records are all-gathered.

`--impl reference` times the UNMODIFIED reference (oracle/_ref/eref_ref) on the host cores, on the headline
workload when the run fits the time budget, else on BASELINE configs[0] — the line says which.
"""
import argparse
import hashlib
import json
import os
import shutil
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "reference Gbp k-mer-screened/sec"
UNIT = "Gbp/s"
WORKLOADS = {
    # name: (genomes per GPU shard, DB bp per shard, query bp total, read length range)
    "c2": (10000, 500_000_000, 50_000_000, (100, 400)),
    "c1": (1000, 50_000_000, 5_000_000, (100, 400)),
    "small": (400, 20_000_000, 2_000_000, (100, 400)),
    "c5": (10000, 500_000_000, 50_000_000, (20, 150)),
}
WORKLOAD_DESC = {
    "c2": "BASELINE configs[1]: 50 Mbp reads vs 10,000-genome ~500 Mbp synthetic phage DB per GPU",
    "c1": "BASELINE configs[0]: 5 Mbp reads vs 1,000-genome ~50 Mbp synthetic phage DB",
    "c5": "BASELINE configs[4]: short-read-heavy 50 Mbp query (reads 20-150 bp) vs 10,000-genome ~500 Mbp synthetic phage DB per GPU",
    "small": "development: 2 Mbp reads vs 400-genome 20 Mbp DB",
}
# strong scaling: ONE database of 8 fixed sub-databases (3,750 genomes, ~187.5 Mbp each), whatever N is
STRONG_SUBS = 8
STRONG = {
    "c3": {"sub": (3750, 187_500_000), "query_bp": 50_000_000,
           "desc": "BASELINE configs[2]: 50 Mbp reads vs ONE 30,000-genome ~1.5 Gbp synthetic phage DB sharded over the GPUs"},
    "c4": {"sub": (3750, 187_500_000), "query_bp": 500_000_000,
           "desc": "BASELINE configs[3]: 500 Mbp reads (~1.3 G query codes) vs ONE 30,000-genome ~1.5 Gbp synthetic phage DB "
                   "sharded over the GPUs"},
}
# algorithmic bytes per unit (restated in DESIGN.md §8)
BYTES_PER_REF_BASE_FULL = 97.0        # 1 B text + 3 probes x 32 B sector
BYTES_PER_QUERY_BASE_BUILD = 193.0    # 1 B text + 3 codes x (32 B sector fill + 32 B write-back)
PLANE_BYTES_PER_BASE = 0.375          # three one-bit planes

_emit = print


def config_for(workload_desc, gpus):
    """the `config` both arms print for the same workload (the driver compares them)"""
    return {"workload": workload_desc, "k": 32, "hit_ratio": 0.9, "perfect_hit_ratio": 0.85, "gpus": gpus}


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks/throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.proc = [], None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", str(index)], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm = [float(r[0]) for r in self.rows if len(r) >= 7 and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) >= 7 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in self.rows if len(r) >= 7 for i in range(4) if r[3 + i].lower().startswith("active")})
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


# ---- synthetic inputs -------------------------------------------------------------------------------

def gen_weak(workload, rank, world, db=None):
    """every rank its own shard (seed 1000 + rank) and 1/world of the reads (seed 2000 + rank), gathered later"""
    from palace_b200 import synth
    G, R, Q, (lo, hi) = WORKLOADS[workload]
    if db is None:
        db = synth.make_db(G, R, 1000 + rank)
    f1, f2, _ = synth.make_reads(db, Q // world, (2000 if workload != "c5" else 5000) + rank, min_len=lo, max_len=hi)
    return db, f1, f2


def gen_strong(kind, rank, world, subs=None):
    """rank r holds sub-databases [8 r/N, 8 (r+1)/N) of the one strong-scaling database and generates the reads drawn
    from them; returns (shard text, fq1 part, fq2 part, bases of the shard, the sub-databases for reuse)"""
    from palace_b200 import synth
    G, R = STRONG[kind]["sub"]
    Q = STRONG[kind]["query_bp"] // STRONG_SUBS
    mine = range(STRONG_SUBS * rank // world, STRONG_SUBS * (rank + 1) // world)
    if subs is None:
        subs = {k: synth.make_db(G, R, 3000 + k) for k in mine}
    f1s, f2s = [], []
    for k in mine:
        a, b, _ = synth.make_reads(subs[k], Q, (4000 if kind == "c3" else 6000) + k)
        f1s.append(a); f2s.append(b)
    text = np.concatenate([subs[k].text for k in mine])
    bases = int(sum(int(subs[k].lengths.sum()) for k in mine))
    return text, np.concatenate(f1s), np.concatenate(f2s), bases, subs


def gather_bytes(t_local, world, device):
    """replicate the query: concatenate every rank's FASTQ part, in rank order (setup, not timed)"""
    import torch
    import torch.distributed as dist
    n = torch.tensor([t_local.numel()], dtype=torch.int64, device=device)
    sizes = [torch.zeros_like(n) for _ in range(world)]
    dist.all_gather(sizes, n)
    sizes = [int(s.item()) for s in sizes]
    slot = max(sizes)
    buf = torch.zeros(slot, dtype=torch.uint8, device=device)
    buf[: t_local.numel()] = t_local
    out = torch.empty(world * slot, dtype=torch.uint8, device=device)
    dist.all_gather_into_tensor(out, buf)
    return torch.cat([out[r * slot: r * slot + sizes[r]] for r in range(world)]), sizes


QUERY_PHASES = ("query_pack_ms", "query_bucket_ms", "query_count_ms")
PHASES = QUERY_PHASES + ("ref_frame_ms", "ref_pack_ms", "scan_summary_ms", "scan_sample_ms", "scan_cand_ms",
                         "scan_probe_ms", "scan_window_ms", "scan_interval_ms")
SCAN_PHASES = tuple(p for p in PHASES if p.startswith("scan_"))
PHASE_STEPS = 3                                # instrumented steps (phase times), not part of the timed region
# kernels of each phase (names as ncu prints them) and how often each runs per step
PHASE_KERNELS = {
    "query_pack_ms": (("k_count_newlines", 2), ("k_fastq_kept", 2), ("k_pack_fastq", 2)),
    "query_bucket_ms": (("k_slice_hist", 2), ("k_slice_offsets", 2), ("k_coarse_scatter_staged", 2), ("k_fine_partition", 2)),
    "query_count_ms": (("k_count_slices_persistent", 1),),
    "ref_frame_ms": (("k_fasta_tiles", 1), ("k_fasta_records", 1), ("k_record_info", 1), ("k_genome_table", 1)),
    "ref_pack_ms": (("k_pack_fasta", 1),),
    "scan_summary_ms": (("k_build_summary", 0),),
    "scan_sample_ms": (("k_probe_sampled", 1),),
    "scan_cand_ms": (("k_need", 1),),
    "scan_probe_ms": (("k_probe_part", 8),),
    "scan_window_ms": (("k_window_list", 1),),
    "scan_interval_ms": (("k_intervals", 1),),
}


def ncu_file():
    for name in ("ncu_kernels_r02.json", "ncu_kernels_r01.json"):
        p = os.path.join(ROOT, "profiles", name)
        if os.path.exists(p):
            return p
    return None


def ncu_traffic(phase):
    """DRAM bytes per step of a phase's kernels from the committed `ncu --set full` capture (profiles/), or None"""
    p = ncu_file()
    if not p:
        return None
    try:
        k = json.load(open(p))["kernels"]
        return int(sum(k[name]["dram_bytes_per_launch"] * times for name, times in PHASE_KERNELS[phase] if times))
    except Exception:
        return None


def records_sha(rec):
    return hashlib.sha256(np.ascontiguousarray(rec).view(np.uint8).tobytes()).hexdigest()[:16]


def prefix_fasta(text, max_records):
    """the first max_records records of a FASTA text (records are independent and ordinals run in file order, so the
    oracle's records of a prefix equal the prefix of the records)"""
    starts = np.flatnonzero(text == 62)
    starts = starts[(starts == 0) | (text[np.maximum(starts, 1) - 1] == 10)]
    if starts.size <= max_records:
        return text
    return text[: int(starts[max_records])]


class Rig:
    """one rank's view of the run: device, stream, process group, the library contexts"""

    def __init__(self, args):
        import torch
        import torch.distributed as dist
        from palace_b200 import eref
        self.torch, self.dist, self.eref = torch, dist, eref
        self.args = args
        self.rank = int(os.environ.get("RANK", "0"))
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        if self.world != args.gpus and self.world > 1:
            raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={self.world}")
        if not torch.cuda.is_available():
            raise SystemExit("bench.py: no CUDA device; the eref path has no CPU fallback")
        torch.cuda.set_device(self.local)
        self.dev = torch.device("cuda", self.local)
        if self.world > 1:
            dist.init_process_group("nccl", device_id=self.dev)
        self.stream = torch.cuda.Stream(device=self.dev)   # the library launches on this stream: its events time the step
        self.coder = eref.random_coder(1)                  # EREF_CODER_SEED=1 convention for GPU-only runs
        # resident_text: the database text of a step is complete in HBM before the step starts, so its framing census
        # may run beside the query work (EREF_FLAG_RESIDENT_TEXT)
        self.g = eref.ErefGPU(self.coder, 0.9, 0.85, device=self.local, prune=not args.no_prune, resident_text=True)
        self.g.set_stream(self.stream.cuda_stream)
        self.g_full = None                                 # second context, candidate filter off (scan_unpruned)
        self.table_mode = "one GPU: whole table built locally"
        self.can_share = False                             # shared reads need the peer pointers
        self.sharing = False
        if self.world > 1:
            self.table_mode = self.wire_table()

    # ---- sharded table build across the ranks ----
    def view(self, ptr, nbytes):
        class _A:
            pass
        a = _A()
        a.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 2}
        return self.torch.as_tensor(a, device=self.dev)

    def all_ok(self, ok):
        t = self.torch.tensor([1 if ok else 0], dtype=self.torch.int32, device=self.dev)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MIN)
        return bool(t.item())

    def wire_table(self):
        g, world, rank = self.g, self.world, self.rank
        if self.args.table == "replicated":
            return "replicated: every GPU counts the whole table"
        if self.args.table in ("auto", "peer"):
            ok = True
            try:
                blob = g.table_export(ipc=True)[0]
            except self.eref.ErefError as e:
                print(f"rank {rank}: no IPC export: {e}", file=sys.stderr)
                ok, blob = False, b""
            blobs = [None] * world
            self.dist.all_gather_object(blobs, blob)
            g.table_set_part(rank, world)
            if ok:
                try:
                    for p in range(world):
                        if p != rank:
                            g.table_set_peer(p, (blobs[p], 0, 0, 0))
                except self.eref.ErefError as e:
                    print(f"rank {rank}: cannot open a peer's table: {e}", file=sys.stderr)
                    ok = False
            if self.all_ok(ok):
                self.can_share = self.args.reads == "shared"
                if self.can_share:
                    return (f"shared reads: each of the {world} GPUs packs and partitions 1/{world} of every query file; a build waits "
                            f"for the peers' lists (device flags), pulls the codes of its table slices out of them over NVLink "
                            f"(CUDA-IPC peer pointers), counts them and stores the finished slices into every GPU's table")
                return (f"shared out over {world} GPUs when the query is large enough (EREF_SHARD_MIN_BYTES, default 160 MB of "
                        f"FASTQ text): each GPU counts its part of the table slices and stores them into every GPU's table "
                        f"through CUDA-IPC peer pointers (NVLink), device-side flags; smaller queries: every GPU counts all")
            if self.args.table == "peer":
                raise SystemExit("--table peer: peer pointers unavailable")
        # hook: one NCCL broadcast per part of the level-3 table and its summary on the library's stream (the parts
        # are not of equal size, so this is not an all-gather in place)
        g.table_set_part(rank, world)                      # (also drops half-wired peers)
        _, bm3, summ, _ = g.table_export()
        full_t = self.view(bm3, 1 << 29)
        full_s = self.view(summ, 1 << 25)

        def hook(_stream):
            for p in range(world):                         # (the parts are re-cut for every query: ask now)
                to, tb, so, sb = g.table_part_range(p)
                self.dist.broadcast(full_t[to: to + tb], src=p)
                self.dist.broadcast(full_s[so: so + sb], src=p)
        g.table_set_hook(hook)
        return (f"shared out over {world} GPUs when the query is large enough (EREF_SHARD_MIN_BYTES): each GPU counts its "
                f"part of the table slices, one NCCL broadcast per part (512 + 32 MiB in all) completes the tables")

    def wire_lists(self, share_bytes):
        """shared reads: a fixed-address arena for this job's lists on every rank, handles exchanged (setup, untimed)"""
        g = self.g
        g.query_set_share(False)
        self.sharing = False
        t = self.torch.tensor([int(share_bytes)], dtype=self.torch.int64, device=self.dev)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        g.query_reserve(int(t.item()) + 4096, 2)
        blobs = [None] * self.world
        self.dist.all_gather_object(blobs, g.lists_export(ipc=True)[0])
        for p in range(self.world):
            if p != self.rank:
                g.lists_set_peer(p, (blobs[p], 0))
        self.barrier()

    def set_sharing(self, on):
        if self.can_share and on != self.sharing:
            self.barrier()
            self.g.query_set_share(on)
            self.sharing = on
            self.barrier()

    def full_ctx(self):
        if self.g_full is None:
            self.g_full = self.eref.ErefGPU(self.coder, 0.9, 0.85, device=self.local, prune=False, resident_text=True)
            self.g_full.set_stream(self.stream.cuda_stream)
        return self.g_full

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def reduce(self, x, op):
        if self.world == 1:
            return x
        t = self.torch.tensor([x], dtype=self.torch.float64, device=self.dev)
        self.dist.all_reduce(t, op={"max": self.dist.ReduceOp.MAX, "sum": self.dist.ReduceOp.SUM,
                                    "min": self.dist.ReduceOp.MIN}[op])
        return float(t.item())

    def device_records(self, ptr, n):
        """zero-copy view of the library's result buffer (n x 4 uint32) as a torch tensor"""
        if n == 0:
            return self.torch.zeros((0, 4), dtype=self.torch.int32, device=self.dev)

        class _Arr:                                        # __cuda_array_interface__ v2
            pass
        a = _Arr()
        a.__cuda_array_interface__ = {"shape": (n, 4), "typestr": "<i4", "data": (ptr, False), "version": 2}
        return self.torch.as_tensor(a, device=self.dev)

    def close(self):
        if self.g_full is not None:
            self.g_full.close()
        self.g.close()
        if self.world > 1:
            self.dist.destroy_process_group()


class Job:
    """one workload on one rank: its texts in HBM and in pinned host memory, and the measurements on it"""

    def __init__(self, rig, fa_np, f1_np, f2_np, shard_bases, label):
        torch = rig.torch
        self.rig, self.label, self.shard_bases = rig, label, shard_bases
        self.info = {}
        rig.barrier()                                      # no rank starts a job's table builds while another still scans
        with torch.cuda.stream(rig.stream):
            self.fa_h = torch.from_numpy(fa_np).pin_memory()
            self.fa_d = self.fa_h.to(rig.dev, non_blocking=True)
            f1_d = torch.from_numpy(f1_np).to(rig.dev)
            f2_d = torch.from_numpy(f2_np).to(rig.dev)
            self.s1_d, self.s2_d = f1_d, f2_d             # this rank's share of the reads (what it generated)
            sizes2 = [int(f2_np.size)]
            if rig.world > 1:                              # the query is the same on every rank
                f1_d, _ = gather_bytes(f1_d, rig.world, rig.dev)
                f2_d, sizes2 = gather_bytes(f2_d, rig.world, rig.dev)
            f1_h, f2_h = f1_d.cpu(), f2_d.cpu()
            # main reads fq2 only up to fq1's size (bin/extract_ref.cpp:1265-1291, eref_fastq_limit): the device-resident
            # steps go through the phase calls, which take a text as it is, so the cut is made here (setup, untimed);
            # the host-buffer call would make it itself.  (No workload of this bench is cut: the files are of one size.)
            lim = rig.eref.fastq_limit(f2_h.numpy(), f1_h.numel())
            if lim < f2_h.numel():
                f2_d, f2_h = f2_d[:lim], f2_h[:lim].clone()
                keep = min(max(lim - sum(sizes2[: rig.rank]), 0), sizes2[rig.rank])
                self.s2_d = self.s2_d[:keep]
            self.info["fq2_bytes_beyond_fq1s_budget"] = int(sum(sizes2)) - int(lim)
            self.f1_d, self.f2_d = f1_d, f2_d
            self.f1_h = f1_h.pin_memory()
            self.f2_h = f2_h.pin_memory()
        rig.stream.synchronize()
        if rig.can_share:
            rig.wire_lists(max(f1_np.size, f2_np.size))
        self.host_cap = int(np.count_nonzero(fa_np == 62)) + 16
        # N > 1, e2e: the query is the same on every rank, so each rank uploads 1/N of it and NCCL all-gathers the
        # text over NVLink; the library takes the gathered (device) text and this rank's database shard (host)
        self.share = None
        if rig.world > 1:
            self.share = []
            for h in (self.f1_h, self.f2_h):
                slot = (h.numel() + rig.world - 1) // rig.world
                lo, hi = min(rig.rank * slot, h.numel()), min((rig.rank + 1) * slot, h.numel())
                self.share.append({"host": h[lo:hi], "mine": torch.zeros(slot, dtype=torch.uint8, device=rig.dev),
                                   "full": torch.empty(rig.world * slot, dtype=torch.uint8, device=rig.dev), "n": h.numel()})

    # ---- steps ----
    def add_query(self, g):
        """both query files: whole (replicated on every rank), or — shared reads — this rank's 1/N of each"""
        if g is self.rig.g and self.rig.sharing:
            g.query_add_device(self.s1_d.data_ptr(), self.s1_d.numel(), want_bytes=False)
            g.query_add_device(self.s2_d.data_ptr(), self.s2_d.numel(), want_bytes=False)
        else:
            g.query_add_device(self.f1_d.data_ptr(), self.f1_d.numel(), want_bytes=False)
            g.query_add_device(self.f2_d.data_ptr(), self.f2_d.numel(), want_bytes=False)

    def step_device(self, g=None, collect=None):
        # no call below waits for its own kernels beyond the sizes the host needs for allocations; the phase times
        # are read (which does wait) only in the instrumented steps, outside the timed region
        from palace_b200.dist import agree_slot, allgather_records
        rig = self.rig
        g = g or rig.g
        g.query_reset()
        self.add_query(g)
        g.ref_set_device(self.fa_d.data_ptr(), self.fa_d.numel())
        ptr, n = g.scan_device(wait=False)
        merged = None
        if rig.world > 1:
            if "slot" not in self.info:
                self.info["slot"] = agree_slot(n, rig.dev)     # once per sharded database
            # (compact=False: the merged records stay on the device with their padding; squeezing it out would read a
            # count back and make the host wait for every step)
            merged = allgather_records(rig.device_records(ptr, n), n, self.info["slot"], compact=False)
        if collect is not None:
            t = g.timings()                                # query times: sums over both files
            for p in PHASES:
                collect[p] = collect.get(p, 0.0) + t[p]
            self.info.update(ref_bases=t["ref_bases"], ref_words=t["ref_words"], need_words=t["ref_need_words"],
                             seq_bytes=t["query_seq_bytes"], n_records=n, query_codes=t["query_codes"])
        return merged

    def gathered_query(self):
        out = []
        for s_ in self.share:
            s_["mine"][: s_["host"].numel()].copy_(s_["host"], non_blocking=True)
            self.rig.dist.all_gather_into_tensor(s_["full"], s_["mine"])
            out.append(self.rig.eref.DeviceBytes(s_["full"].data_ptr(), s_["n"]))
        return out

    def step_host(self):
        """the public host-buffer call: pinned host -> (overlapped H2D) -> kernels -> D2H of the records"""
        from palace_b200.dist import allgather_records
        rig = self.rig
        if rig.world > 1:
            q1, q2 = self.gathered_query()
            rec = rig.g.run_host(q1, q2, self.fa_h.numpy(), chunks=rig.args.chunks, cap=self.host_cap)
            ptr, n = rig.g.run_results_device()            # the records are still in HBM: all-gather them from there
            merged = allgather_records(rig.device_records(ptr, n), n, self.info.get("slot"))
            self.last_host_records = rec
            return merged.cpu() if rig.rank == 0 else None
        rec = rig.g.run_host(self.f1_h.numpy(), self.f2_h.numpy(), self.fa_h.numpy(), chunks=rig.args.chunks, cap=self.host_cap)
        self.last_host_records = rec
        return rec

    # ---- measurements ----
    def time_device(self, steps, warmup):
        rig, torch = self.rig, self.rig.torch
        rig.set_sharing(True)
        acc = {}
        with torch.cuda.stream(rig.stream):
            for _ in range(warmup):
                self.step_device()
            # instrumented steps, outside the timed region: the library's streams serialised (eref_set_serial) so that every
            # phase time is the duration of its own kernels; the timed steps overlap them and are shorter than the sum
            rig.g.set_serial(True)
            for _ in range(PHASE_STEPS):
                self.step_device(collect=acc)
            rig.g.set_serial(False)
            self.step_device()
            rig.barrier()
            launches0 = rig.g.timings()["kernel_launches"]
            sampler = ClockSampler(rig.local) if rig.rank == 0 and not os.environ.get("EREF_BENCH_NO_SAMPLER") else None
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            marks = [torch.cuda.Event(enable_timing=True) for _ in range(steps)]      # step-by-step spread (diagnostic)
            e0.record(rig.stream)
            for k in range(steps):
                self.step_device()
                marks[k].record(rig.stream)
            e1.record(rig.stream)
            rig.barrier()
            dev_ms = e0.elapsed_time(e1)
            per = [([e0] + marks)[k].elapsed_time(marks[k]) for k in range(steps)]
            self.step_spread = {"min": round(min(per), 4), "median": round(float(np.median(per)), 4), "max": round(max(per), 4)}
            launches = rig.g.timings()["kernel_launches"] - launches0
            clocks = sampler.stop() if sampler else None
        self.avg = {p: acc[p] / PHASE_STEPS for p in PHASES}
        if rig.sharing:                                    # every rank saw its share of the reads only
            self.info["seq_bytes"] = int(rig.reduce(float(self.info["seq_bytes"]), "sum"))
            self.info["query_codes"] = int(rig.reduce(float(self.info["query_codes"]), "sum"))
        if rig.world > 1:
            to, tb, _, _ = rig.g.table_part_range()
            self.info["table_build"] = ("every GPU counts the whole table (query below EREF_SHARD_MIN_BYTES)" if tb == 1 << 29 else
                                        f"shared out: rank 0 counts table bytes [{to}, {to + tb}) of 2^29 and receives the rest")
        ms = rig.reduce(dev_ms, "max") / steps
        self.total_bases = rig.reduce(float(self.info["ref_bases"]), "sum")
        return {"ms_per_step": ms, "value": self.total_bases / 1e9 / (ms / 1e3),
                "launches": int(rig.reduce(float(launches), "sum")), "clocks": clocks, "step_spread_rank0": self.step_spread}

    def time_host(self, steps, warmup):
        rig, torch = self.rig, self.rig.torch
        rig.set_sharing(False)                             # eref_run_host takes whole files
        with torch.cuda.stream(rig.stream):
            for _ in range(min(warmup, 2)):
                self.step_host()
            rig.barrier()
            t0 = time.perf_counter()
            for _ in range(steps):
                self.step_host()
            rig.barrier()
            e2e_s = time.perf_counter() - t0
            # what the bus alone takes: the same pinned buffers copied to the device, nothing else running
            c0, c1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            srcs = (self.f1_h, self.f2_h, self.fa_h) if rig.world == 1 else (self.share[0]["host"], self.share[1]["host"], self.fa_h)
            h2d_bytes = sum(int(x.numel()) for x in srcs)
            scratch = [torch.empty(x.numel(), dtype=torch.uint8, device=rig.dev) for x in srcs]
            for _ in range(2):
                c0.record(rig.stream)
                for dst, src in zip(scratch, srcs):
                    dst.copy_(src, non_blocking=True)
                c1.record(rig.stream)
                rig.stream.synchronize()
            h2d_ms = c0.elapsed_time(c1)
            # N > 1: the same copies with the ranks taking turns (every other GPU idle) — against the concurrent figure
            # above this tells how much of an end-to-end slowdown is the host side of PCIe shared by the ranks
            solo_ms = None
            if rig.world > 1:
                for r in range(rig.world):
                    rig.barrier()
                    if r == rig.rank:
                        c0.record(rig.stream)
                        for dst, src in zip(scratch, srcs):
                            dst.copy_(src, non_blocking=True)
                        c1.record(rig.stream)
                        rig.stream.synchronize()
                        solo_ms = c0.elapsed_time(c1)
                rig.barrier()
            del scratch
        e2e_s = rig.reduce(e2e_s, "max")
        h2d_ms = rig.reduce(h2d_ms, "max")
        solo = None
        if solo_ms is not None:
            solo = {"h2d_solo_GBps_min": round(h2d_bytes / 1e6 / rig.reduce(solo_ms, "max"), 1),
                    "h2d_concurrent_GBps_min": round(h2d_bytes / 1e6 / h2d_ms, 1),
                    "note": "per-GPU host->device rate of this step's buffers: one rank at a time vs all ranks at once"}
        return {"value": round(self.total_bases / 1e9 / (e2e_s / steps), 3), "unit": UNIT,
                "h2d_bytes_per_step": h2d_bytes, "d2h_bytes_per_step": int(self.info["n_records"]) * 16 + 24,
                "ms_per_step": round(e2e_s / steps * 1e3, 3),
                "api": f"eref_run_host(chunks={rig.args.chunks}): pinned host buffers, H2D overlapped with kernels" +
                       ("" if rig.world == 1 else "; each rank uploads 1/N of the query text, NCCL all-gathers it over NVLink"),
                "h2d_alone_ms": round(h2d_ms, 3), "h2d_alone_GBps": round(h2d_bytes / 1e6 / h2d_ms, 1), "h2d_sharing": solo}

    def time_resident(self, steps):
        """NOT headline numbers: the same sample against a database whose planes stay in HBM (eref_ref_set once)"""
        from palace_b200.dist import allgather_records
        rig, torch = self.rig, self.rig.torch
        g = rig.g

        def step_resident():
            g.query_reset()
            self.add_query(g)
            ptr, n = g.scan_device(wait=False)
            if rig.world > 1:
                allgather_records(rig.device_records(ptr, n), n, self.info["slot"], compact=False)

        def step_host_resident():
            if rig.world > 1:
                q1, q2 = self.gathered_query()
                rec = g.run_host(q1, q2, None, cap=self.host_cap)
                t = torch.from_numpy(rec.view(np.uint32).reshape(-1, 4).astype(np.int32)).to(rig.dev)
                allgather_records(t, rec.size, self.info.get("slot"))
            else:
                g.run_host(self.f1_h.numpy(), self.f2_h.numpy(), None, cap=self.host_cap)
        rig.set_sharing(True)
        with torch.cuda.stream(rig.stream):
            g.ref_set_device(self.fa_d.data_ptr(), self.fa_d.numel())
            step_resident()
            rig.barrier()
            r0, r1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            r0.record(rig.stream)
            for _ in range(steps):
                step_resident()
            r1.record(rig.stream)
            rig.barrier()
            res_ms = rig.reduce(r0.elapsed_time(r1), "max") / steps
        rig.set_sharing(False)
        with torch.cuda.stream(rig.stream):
            step_host_resident()
            rig.barrier()
            t0 = time.perf_counter()
            for _ in range(steps):
                step_host_resident()
            rig.barrier()
            e2e_res_s = rig.reduce(time.perf_counter() - t0, "max") / steps
        return ({"ms_per_step": round(res_ms, 4), "value": round(self.total_bases / 1e9 / (res_ms / 1e3), 2), "unit": UNIT,
                 "note": "not the headline: the same step when the database's bit-planes stay in HBM across samples "
                         "(eref_ref_set once), i.e. table build + scan + all-gather"},
                {"ms_per_step": round(e2e_res_s * 1e3, 3), "value": round(self.total_bases / 1e9 / e2e_res_s, 2), "unit": UNIT,
                 "h2d_bytes_per_step": int(self.f1_h.numel() + self.f2_h.numel()),
                 "note": "not the headline: eref_run_host(fq1, fq2, NULL) from pinned host buffers against a database set "
                         "once (one phage DB, many samples)"})

    def scan_unpruned(self, steps=3):
        """The scan with the candidate filter OFF (EREF_FLAG_NO_PRUNE): all three probes at every reference position, timed
        by the library's CUDA events in this run — the figure comparable with SURVEY 8(d)'s 97 B per reference base
        (1 B of text + 3 probes x one 32 B DRAM sector).  The table is built locally by this second context."""
        rig, torch = self.rig, self.rig.torch
        if rig.args.no_prune:
            g, own = rig.g, False
        else:
            g, own = rig.full_ctx(), True
        peak, _ = peaks()
        ms = []
        with torch.cuda.stream(rig.stream):
            g.query_reset()
            if g is rig.g:
                rig.set_sharing(False)
            self.add_query(g)
            g.ref_set_device(self.fa_d.data_ptr(), self.fa_d.numel())
            for i in range(steps + 1):
                ptr, n = g.scan_device(wait=True)
                t = g.timings()
                if i:
                    ms.append(sum(t[p] for p in SCAN_PHASES))
            rec_full = g.scan().copy() if own else None
        t_ms = rig.reduce(float(np.mean(ms)), "max")
        R = rig.reduce(float(t["ref_bases"]), "sum")
        out = {"ms": round(t_ms, 4), "steps": steps, "value": round(R / 1e9 / (t_ms / 1e3), 2), "unit": UNIT,
               "bytes_per_ref_base": BYTES_PER_REF_BASE_FULL,
               "frac": round(BYTES_PER_REF_BASE_FULL * R / 1e9 / (t_ms / 1e3) / (peak * rig.world), 4),
               "probes_per_s_G": round(3.0 * R / (t_ms / 1e3) / 1e9, 1),
               "phases_ms": {p: round(t[p], 4) for p in SCAN_PHASES if t[p] > 0},
               "note": "frac = 97 B x reference bases / scan time / (N x measured HBM peak); above 1 would mean the "
                       "L2-sized table parts answer probes that the 97 B model charges to DRAM"}
        return out, rec_full

    def parity(self, rec_gpu, max_records=None):
        """The records of the timed workload against the CPU oracle on the same texts (test infrastructure, outside every
        timed region).  max_records: compare a prefix of the shard's records (see prefix_fasta).  At N > 1 rank 0 checks
        its own shard with all host cores (every rank would have to count the whole query again)."""
        from tests import oracle_util
        rig = self.rig
        if rig.rank != 0:
            return None
        orc = oracle_util.load()
        fa = self.fa_h.numpy()
        if max_records is not None:
            fa = prefix_fasta(fa, max_records)
        threads = os.cpu_count() or 1
        t0 = time.perf_counter()
        want = orc.run(self.f1_h.numpy(), self.f2_h.numpy(), fa, rig.coder, 0.9, 0.85, threads=threads)
        dt = time.perf_counter() - t0
        got = np.ascontiguousarray(rec_gpu[: want.size])
        same = bool(got.size == want.size and (got == want).all())
        sel = int(np.count_nonzero((want["el"] > 0) & (want["el"].astype(np.float32) / want["ref_len"].astype(np.float32) > 0.75)))
        out = {"checked": True, "records": int(want.size), "selected": sel, "identical": same,
               "of_records_on_rank0": int(len(rec_gpu)),
               "against": f"oracle port (oracle/liberef_oracle.so, {threads} threads), {dt:.0f} s"
                          + ("" if rig.world == 1 else "; rank 0's shard"),
               "sha256_16": records_sha(got)}
        if not same:
            bad = np.flatnonzero(~(got == want[: got.size])) if got.size == want.size else []
            out["first_mismatch"] = [int(x) for x in bad[:4]]
        return out


def phases_report(job, fa_bytes, ntext_q, no_prune):
    """per-phase times with (i) SURVEY 8(d)'s algorithmic bytes where it gives a figure, (ii) the bytes this design must
    move, (iii) ncu's DRAM bytes from the committed capture"""
    peak, _ = peaks()
    info, avg = job.info, job.avg
    qbases, codes, R = info["seq_bytes"], info["query_codes"], info["ref_bases"]
    probed = R if no_prune else 32 * info["need_words"]
    bytes_8d = {
        "query_pack_ms": ntext_q + PLANE_BYTES_PER_BASE * qbases,            # K1: text once, planes once
        "ref_pack_ms": fa_bytes + PLANE_BYTES_PER_BASE * R,                  # K1: text once, planes once
        "ref_frame_ms": None,                                                # framing is part of K1's single text read
        "scan_probe_ms": BYTES_PER_REF_BASE_FULL * probed,                   # 1 B + 3 x 32 B per probed base
    }
    design = {
        "query_pack_ms": 3.0 * ntext_q + PLANE_BYTES_PER_BASE * qbases,      # census, kept, pack read the text; planes written
        "query_bucket_ms": 2 * PLANE_BYTES_PER_BASE * qbases + 12.0 * codes, # planes read twice; codes written, read, written
        "query_count_ms": 4.0 * codes + 4.0 * (1 << 27) + 4.0 * (1 << 23),   # codes read; level 3 + summary written
        "ref_frame_ms": 1.0 * fa_bytes,
        "ref_pack_ms": 1.0 * fa_bytes + 2 * (PLANE_BYTES_PER_BASE + 0.125) * R,   # planes + record ids zeroed, then written
        "scan_sample_ms": (PLANE_BYTES_PER_BASE + 1.0 / 32) * R,             # planes read, hit bytes written; probes: L2 + ncu
        "scan_cand_ms": (1 + 4.0 / 8 + 1 + 12) / 32.0 * R,                   # hits, record ids, need, zeroed single/trio/good
        "scan_probe_ms": (96.0 + 0.625) * probed,
        "scan_window_ms": (0.25 + 0.125) * probed,
        "scan_interval_ms": 0.125 * R,
    }
    out = {}
    for p in PHASES:
        e = {"ms": round(avg[p], 4), "kernels": [n for n, t in PHASE_KERNELS[p] if t],
             "design_bytes": int(design.get(p, 0)), "ncu_dram_bytes": ncu_traffic(p)}
        if avg[p] > 0:
            if bytes_8d.get(p):
                e["survey_8d_bytes"] = int(bytes_8d[p])
                e["frac_of_hbm"] = round(bytes_8d[p] / 1e9 / (avg[p] / 1e3) / peak, 3)
            e["design_frac_of_hbm"] = round(design.get(p, 0) / 1e9 / (avg[p] / 1e3) / peak, 3)
        out[p] = e
    return out, bytes_8d, design, probed


def roofline_record(job, phases, bytes_8d, design, probed, no_prune):
    """the dominant KERNEL = the longest of the phases that consist of one kernel; achieved = SURVEY 8(d) algorithmic
    bytes / its CUDA-event time where 8(d) gives a figure (pack: text once + planes once; probe: 97 B per probed base),
    else the bytes the kernel must move (count: codes read + table written; sampled probe: ncu's DRAM bytes)"""
    peak, peak_src = peaks()
    avg, info = job.avg, job.info
    single = {"scan_probe_ms": "k_probe_full" if no_prune else "k_probe_part (8 launches: the table in 8 parts)",
              "scan_sample_ms": "k_probe_sampled", "ref_pack_ms": "k_pack_fasta", "query_count_ms": "k_count_slices_persistent"}
    dom = max(single, key=lambda p_: avg[p_])
    R, codes = info["ref_bases"], info["query_codes"]
    if dom == "scan_probe_ms":
        alg, src, units, unit_name = bytes_8d[dom], "SURVEY 8(d): 97 B per reference base probed", probed, \
            "reference bases probed by the launches (all three streams)"
        note = ("random 4-byte probes; walking the table in L2-sized parts turns most into L2 hits, so ncu's DRAM traffic "
                "is below the algorithmic bytes and the rounds are instruction-bound")
    elif dom == "ref_pack_ms":
        alg, src, units, unit_name = bytes_8d[dom], "SURVEY 8(d) K1: text read once + planes written once", R, "reference bases"
        note = "instruction-bound text kernel (integer SIMD-in-register masks); design_frac also counts the plane memsets"
    elif dom == "query_count_ms":
        alg, src, units, unit_name = design[dom], "codes read once (4 B) + level-3 table and summary written once", codes, \
            "query codes (3 per valid 32-mer)"
        note = "bound by shared-memory atomics, one CTA per SM"
    else:
        t = ncu_traffic(dom)
        alg = t if t else design[dom]
        src = "bytes actually moved (ncu DRAM bytes of the committed capture)" if t else "design bytes (planes + hit bytes)"
        units, unit_name = R, "reference bases"
        note = "bound by one scattered L2 sector per clock and SM (most sampled probes are answered by the 32 MiB summary in L2)"
    achieved = alg / 1e9 / (avg[dom] / 1e3) if avg[dom] > 0 else 0.0
    traffic = ncu_traffic(dom)
    return {"bound": "hbm", "kernel": single[dom], "phase": dom, "achieved": round(achieved, 1), "peak": peak, "unit": "GB/s",
            "frac": round(achieved / peak, 4), "traffic": traffic,
            "traffic_source": (os.path.relpath(ncu_file(), ROOT) + " (committed ncu --set full capture of this command, "
                               "not measured in this run)") if traffic else None,
            "peak_source": peak_src, "bytes_source": src,
            "algorithmic_bytes_per_unit": round(alg / max(units, 1), 3), "units_per_launch": int(units), "unit_name": unit_name,
            "algorithmic_bytes_per_launch": int(alg), "kernel_ms": round(avg[dom], 4),
            "design_frac": round(design[dom] / 1e9 / (avg[dom] / 1e3) / peak, 4) if avg[dom] > 0 else None,
            "probes_per_s_G": round(3.0 * probed / (avg["scan_probe_ms"] / 1e3) / 1e9, 1) if avg["scan_probe_ms"] > 0 else None,
            "note": note, "largest_phase": max(avg, key=avg.get)}


def sub_record(rig, job, steps, warmup, parity_records, with_e2e=True, with_unpruned=True):
    """device-resident time, e2e, unpruned scan and parity of one extra workload (strong / c5)"""
    t = job.time_device(steps, warmup)
    out = {"ms_per_step": round(t["ms_per_step"], 4), "value": round(t["value"], 3), "unit": UNIT,
           "db_bases_total": int(job.total_bases), "db_bases_this_gpu": int(job.info["ref_bases"]),
           "query_seq_bases": int(job.info["seq_bytes"]), "query_codes": int(job.info["query_codes"]),
           "phases_ms": {k: round(v, 4) for k, v in job.avg.items()},
           "need_word_fraction": round(job.info["need_words"] / max(job.info["ref_words"], 1), 4),
           "gpu_launches": t["launches"], "step_spread_ms_rank0": t["step_spread_rank0"]}
    if "table_build" in job.info:
        out["table_build"] = job.info["table_build"]
    scan_ms = sum(job.avg[p] for p in SCAN_PHASES)
    out["scan_only"] = {"ms": round(scan_ms, 4), "value": round(job.info["ref_bases"] * rig.world / 1e9 / (scan_ms / 1e3), 2), "unit": UNIT}
    if with_e2e:
        e = job.time_host(max(3, steps // 2), 2)
        out["e2e"] = {k: e[k] for k in ("value", "unit", "ms_per_step", "h2d_bytes_per_step", "d2h_bytes_per_step")}
    if with_unpruned:
        out["scan_unpruned"], _ = job.scan_unpruned()
    if not rig.args.no_parity:
        rig.set_sharing(False)
        with rig.torch.cuda.stream(rig.stream):
            job.step_host()                                # (a collective at N > 1: every rank takes part)
        out["parity"] = job.parity(job.last_host_records.copy(), parity_records)
    return out


def run_b200(args):
    t_begin = time.perf_counter()
    rig = Rig(args)
    rank, world = rig.rank, rig.world
    peak, _ = peaks()
    strong_kinds = [k for k in args.strong.split(",") if k and k != "none"]
    for k in strong_kinds:
        if k not in STRONG:
            raise SystemExit(f"--strong {k}: not one of {sorted(STRONG)}")
    if world not in (1, 2, 4, 8):
        strong_kinds = []

    # ---- headline: weak-scaled BASELINE configs[1] (or --workload) ----
    db, f1, f2 = gen_weak(args.workload, rank, world)
    job = Job(rig, db.text, f1, f2, int(db.lengths.sum()), args.workload)
    head = job.time_device(args.steps, args.warmup)
    db_res, e2e_res = job.time_resident(args.steps)
    e2e = job.time_host(args.steps, args.warmup)
    unpruned, rec_full = None, None
    if not args.no_unpruned:
        unpruned, rec_full = job.scan_unpruned()
    par = None
    if not args.no_parity:
        rig.set_sharing(False)
        with rig.torch.cuda.stream(rig.stream):
            job.step_host()
            rec_host = job.last_host_records.copy()
        # and the device-resident path that `value` times (at N > 1: shared reads), against the host-buffer call's records
        rig.set_sharing(True)
        with rig.torch.cuda.stream(rig.stream):
            g = rig.g
            g.query_reset()
            job.add_query(g)
            g.ref_set_device(job.fa_d.data_ptr(), job.fa_d.numel())
            rec_dev = g.scan().copy()
        rig.set_sharing(False)
        par = job.parity(rec_host)
        same_paths = rig.reduce(1.0 if (rec_dev.size == rec_host.size and (rec_dev == rec_host).all()) else 0.0, "min") > 0.5
        same_full = rig.reduce(1.0 if (rec_full is None or (rec_full.size == rec_host.size and (rec_full == rec_host).all()))
                               else 0.0, "min") > 0.5
        if par is not None:
            par["device_path_equals_host_path"] = bool(same_paths)
            if rec_full is not None:
                par["unpruned_equals_pruned"] = bool(same_full)
    phases, b8d, design, probed = phases_report(job, job.fa_d.numel(), job.f1_d.numel() + job.f2_d.numel(), args.no_prune)
    roof = roofline_record(job, phases, b8d, design, probed, args.no_prune)
    info, avg = dict(job.info), dict(job.avg)
    qbases = info["seq_bytes"]
    scan_ms = sum(avg[p] for p in SCAN_PHASES)
    head_fa_bytes, head_q_bytes = job.fa_d.numel(), job.f1_d.numel() + job.f2_d.numel()

    # ---- c5: short-read-heavy query against the same database ----
    c5 = None
    if args.workload == "c2" and not args.no_c5:
        _, s1, s2 = gen_weak("c5", rank, world, db=db)
        job5 = Job(rig, db.text, s1, s2, int(db.lengths.sum()), "c5")
        c5 = sub_record(rig, job5, max(5, args.steps // 2), 3, parity_records=None, with_e2e=True, with_unpruned=False)
        c5["workload"] = WORKLOAD_DESC["c5"]
        c5["reads"] = "20-150 bp (40 % shorter than k = 32 contribute nothing, bin/extract_ref.cpp:964); k sweep = the single point k = 32 (:24)"
        del job5
    del job, db, f1, f2
    rig.torch.cuda.empty_cache()

    # ---- strong scaling: ONE 30,000-genome ~1.5 Gbp database over the N GPUs ----
    strong = {}
    subs = None
    for kind in strong_kinds:
        text, q1, q2, bases, subs = gen_strong(kind, rank, world, subs)
        js = Job(rig, text, q1, q2, bases, kind)
        del text, q1, q2
        rec = sub_record(rig, js, max(5, args.steps // 2), 3, parity_records=1500 if kind == "c4" else 7500,
                         with_e2e=True, with_unpruned=True)
        rec["workload"] = STRONG[kind]["desc"]
        rec["scaling"] = "strong"
        rec["n_gpus"] = world
        strong[kind] = rec
        del js
        rig.torch.cuda.empty_cache()
    del subs

    if rank != 0:
        rig.close()
        return
    out = {
        "metric": METRIC, "value": round(head["value"], 3), "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": round(head["ms_per_step"], 4), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u32", "data": "synthetic",
        "config": config_for(WORKLOAD_DESC[args.workload], world),
        "workload_detail": {"db_bases_per_gpu": int(info["ref_bases"]), "query_seq_bases": int(qbases),
                            "records_per_gpu": int(info["n_records"]),
                            "parallelism": f"db-shard x{world}, query replicated", "table_build": rig.table_mode,
                            "table_build_this_query": info.get("table_build"),
                            "candidate_filter": not args.no_prune,
                            "l2": "inputs larger than L2: per step the 512 MiB level-3 table is rewritten and 0.6 GB of text read"},
        "e2e": e2e,
        "gpu_launches": head["launches"],
        "clocks": head["clocks"],
        "step_spread_ms_rank0": head["step_spread_rank0"],
        "phases_ms": {k: round(v, 4) for k, v in avg.items()},
        "scan_only": {"ms": round(scan_ms, 4), "value": round(info["ref_bases"] * world / 1e9 / (scan_ms / 1e3), 2), "unit": UNIT,
                      "need_word_fraction": round(info["need_words"] / max(info["ref_words"], 1), 4),
                      "note": "pruned scan: the exact candidate filter skips the probes of words no good window can use, so "
                              "this is NOT comparable with 97 B/base; scan_unpruned is"},
        "scan_unpruned": unpruned,
        "parity": par,
        "strong": strong or None,
        "c5": c5,
        "db_resident": db_res, "e2e_db_resident": e2e_res,
        "roofline": roof,
        "phases": phases,
        "survey_8d": {"table_build": {"bytes_per_query_base": BYTES_PER_QUERY_BASE_BUILD,
                                      "ms": round(sum(avg[p] for p in QUERY_PHASES), 4),
                                      "frac_of_hbm": round(BYTES_PER_QUERY_BASE_BUILD * qbases / 1e9 /
                                                           (sum(avg[p] for p in QUERY_PHASES) / 1e3) / peak, 3),
                                      "note": "193 B per query base assumes three random DRAM read-modify-writes; the "
                                              "partitioned shared-memory build replaces them"},
                      "scan": "see scan_unpruned (same run, candidate filter off)"},
    }
    if world == 1 and not args.no_cpu_baseline:
        out["cpu_baseline"] = cpu_baseline(budget_runs=1)
    if world > 1 and not args.no_exe_check:
        out["exe_multi_gpu"] = exe_multi_gpu(world)
    out["bench_wall_s"] = round(time.perf_counter() - t_begin, 1)     # everything this command did, setup and checks included
    _emit(json.dumps(out))
    rig.close()


# ---- the reference's CPU implementation on the host cores -----------------------------------------

REF_BIN = os.path.join(ROOT, "oracle", "_ref", "eref_ref")
ORACLE_BIN = os.path.join(ROOT, "oracle", "eref_oracle")


def _host_mem_gb():
    try:
        for line in open("/proc/meminfo"):
            if line.startswith("MemAvailable"):
                return int(line.split()[1]) / 1e6
    except Exception:
        pass
    return 0.0


def _disk_free_gb(path):
    try:
        return shutil.disk_usage(path).free / 1e9
    except Exception:
        return 0.0


def have_reference():
    return os.path.exists(REF_BIN) and os.access(REF_BIN, os.X_OK) and _host_mem_gb() > 30


class CpuArm:
    """One workload of the bench generator on disk, and one timed invocation of the reference's CPU implementation per
    step.  kind = "reference": the UNMODIFIED reference binary built from /root/reference/bin/extract_ref.cpp (needs
    ~24 GB RAM, seconds of fixed cost per run, a 12 B/base index next to the database); kind = "port": oracle/eref_oracle
    (the C restatement) only when the binary is missing or the host has too little memory for it."""

    def __init__(self, sample="c1", kind=None):
        from palace_b200 import synth
        G, R, Q, (lo, hi) = WORKLOADS[sample]
        self.sample_name = sample
        self.dir = tempfile.mkdtemp(prefix="eref_cpu_")
        db = synth.make_db(G, R, 1000)
        f1, f2, _ = synth.make_reads(db, Q, 2000, min_len=lo, max_len=hi)
        db.text.tofile(os.path.join(self.dir, "db.fa"))
        f1.tofile(os.path.join(self.dir, "r1.fq")); f2.tofile(os.path.join(self.dir, "r2.fq"))
        self.bases = int(db.lengths.sum())
        self.cores = os.cpu_count() or 1
        if kind is None:
            kind = "reference" if have_reference() else "port"
        self.kind = kind
        self.bin = REF_BIN if kind == "reference" else ORACLE_BIN
        if kind == "port" and not os.path.exists(ORACLE_BIN):
            subprocess.run(["make", "-C", os.path.join(ROOT, "oracle"), "eref_oracle"], check=True, capture_output=True)
        self.sample = (f"{WORKLOAD_DESC[sample]}: {G}-genome {self.bases / 1e6:.0f} Mbp DB + {Q / 1e6:.0f} Mbp reads (same "
                       f"generator and seeds as the GPU arm's rank 0), threads={self.cores}, "
                       + ("unmodified reference binary oracle/_ref/eref_ref" if kind == "reference" else "oracle port oracle/eref_oracle"))
        self.stdout = None

    def run_once(self, threads=None, fq2="r2.fq", timeout=None):
        env = dict(os.environ, EREF_CODER_SEED="1", EREF_ORACLE_WRITE_INDEX="1")
        t0 = time.perf_counter()
        r = subprocess.run([self.bin, "r1.fq", fq2, "db.fa", "tmp.txt", "0.9", "0.85", str(threads or self.cores)],
                           cwd=self.dir, capture_output=True, env=env, timeout=timeout)
        dt = time.perf_counter() - t0
        if r.returncode != 0:
            raise RuntimeError(f"{self.bin} failed: {r.stderr.decode()[-400:]}")
        self.stdout = r.stdout
        return dt

    def run_product(self, gpus=1, fq2="r2.fq"):
        """palace_b200/bin/eref as a process on the very same files (it takes the coder from the index header the
        CPU run left next to the database): wall time and stdout"""
        exe = os.path.join(ROOT, "palace_b200", "bin", "eref")
        t0 = time.perf_counter()
        r = subprocess.run([exe, "r1.fq", fq2, "db.fa", "tmp_b200.txt", "0.9", "0.85", str(self.cores)], cwd=self.dir,
                           capture_output=True, env=dict(os.environ, EREF_GPUS=str(gpus)))
        dt = time.perf_counter() - t0
        if r.returncode != 0:
            raise RuntimeError(f"{exe} failed: {r.stderr.decode()[-400:]}")
        return dt, r.stdout

    def close(self):
        shutil.rmtree(self.dir, ignore_errors=True)


def cpu_baseline(budget_runs=1):
    """N=1 line: the reference on the host cores on a BOUNDED sample (BASELINE configs[0] size), and the whole-process
    comparison on that sample's files"""
    arm = CpuArm("c1")
    try:
        cold = arm.run_once()                          # builds <db>.k32.index.dat serially (:1250)
        warm = min(arm.run_once() for _ in range(budget_runs))
        out = {"value": round(arm.bases / 1e9 / warm, 6), "unit": UNIT, "cores": arm.cores, "kind": arm.kind,
               "sample": arm.sample + "; value = warm run (index on disk)", "sample_db_bases": arm.bases,
               "warm_s": round(warm, 2), "cold_s": round(cold, 2), "cold_value": round(arm.bases / 1e9 / cold, 6),
               "selected_refs": arm.stdout.count(b"\n")}
        # wall time of the whole eref PROCESS, both executables on the same files of this sample, and
        # the same-run parity check: the reference at threads=1 (the parity configuration) against the drop-in's stdout
        try:
            t1 = arm.run_once(threads=1)
            ref_t1_stdout = arm.stdout
            first, _ = arm.run_product()                  # includes a cold CUDA context and module load
            again, mine = arm.run_product()
            out["process"] = {"reference_threads1_wall_s": round(t1, 2), "reference_wall_s": round(warm, 2),
                              "b200_wall_s": round(min(first, again), 3), "b200_first_run_wall_s": round(first, 3),
                              "stdout_identical_to_reference_threads1": mine == ref_t1_stdout,
                              "result_lines": mine.count(b"\n"),
                              "note": "one process per sample as palace runs it: file reads, CUDA context, upload, kernels, output"}
        except Exception as e:                            # the bench line must not depend on this extra
            out["process"] = {"error": str(e)[-300:]}
        # ... and once more with fq2 the LONGER file (its own text followed by the first third of fq1's): main hands
        # file_size(fq1) to read_fastq as the byte budget of both files (bin/extract_ref.cpp:1265-1291, :942-945), so the
        # reference never reads the tail; the drop-in must print the same lines
        try:
            if "error" not in out["process"]:
                with open(os.path.join(arm.dir, "r1.fq"), "rb") as f:
                    r1 = f.read()
                with open(os.path.join(arm.dir, "r2.fq"), "rb") as f:
                    r2 = f.read()
                with open(os.path.join(arm.dir, "r2_longer.fq"), "wb") as f:
                    f.write(r2 + r1[: len(r1) // 3])
                arm.run_once(threads=1, fq2="r2_longer.fq")
                ref_longer = arm.stdout
                _, mine_longer = arm.run_product(fq2="r2_longer.fq")
                out["process"]["fq2_longer_than_fq1"] = {
                    "stdout_identical_to_reference_threads1": mine_longer == ref_longer,
                    "result_lines": mine_longer.count(b"\n"),
                    "reference_stdout_equals_the_plain_files": ref_longer == ref_t1_stdout,
                    "note": "the reference reads fq2 only up to fq1's size: all but the first ~1 % of the appended reads are never counted"}
        except Exception as e:
            out["process"]["fq2_longer_than_fq1"] = {"error": str(e)[-300:]}
        return out
    finally:
        arm.close()


def exe_multi_gpu(world):
    """N>1 line, rank 0: the drop-in EXECUTABLE with EREF_GPUS=N (one host thread per GPU, sharded table build, NCCL
    all-gather of the records) next to the unmodified reference at threads=1 on the configs[0]-sized files"""
    try:
        arm = CpuArm("c1")
    except Exception as e:
        return {"error": str(e)[-300:]}
    try:
        arm.run_once()                                 # writes the index (coder header) the drop-in reads
        t1 = arm.run_once(threads=1) if arm.kind == "reference" else arm.run_once()
        ref_stdout = arm.stdout
        wall, mine = arm.run_product(gpus=world)
        one, mine1 = arm.run_product(gpus=1)
        return {"n": world, "stdout_identical": mine == ref_stdout, "stdout_identical_1gpu": mine1 == ref_stdout,
                "wall_s": round(wall, 3), "wall_s_1gpu": round(one, 3), "reference_threads1_wall_s": round(t1, 2),
                "result_lines": mine.count(b"\n"), "against": arm.kind,
                "note": "the other ranks of this bench hold their GPUs meanwhile"}
    except Exception as e:
        return {"n": world, "error": str(e)[-300:]}
    finally:
        arm.close()


def run_reference(args):
    """The reference's own CPU implementation of the path on the host cores, all threads: the UNMODIFIED reference binary
    whenever it exists, each step one whole eref process with the index already on disk (the index build is untimed
    setup, reported as cold_s).  It runs the headline workload when the whole arm can end within EREF_REF_LIMIT_S
    (default 780 s: the driver cuts a command of its scaling run at 870 s), else BASELINE configs[0]; config.workload,
    `sample` and `chosen_because` name what actually ran and why.  configs[0] runs first (with its index on disk it is the
    fallback, and its time per process says how much the fallback needs); then the headline workload is tried: its index
    build is abandoned if going on would leave the fallback no time, and after its first warm process the whole arm is
    projected once more."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    t_begin = time.perf_counter()
    limit = float(os.environ.get("EREF_REF_LIMIT_S", "780"))
    forced = os.environ.get("EREF_BENCH_CPU_SAMPLE")          # the CPU test-suite: "small"
    kind = os.environ.get("EREF_BENCH_CPU_KIND") or None      # ... and "port"
    runs = args.steps + args.warmup
    elapsed = lambda: time.perf_counter() - t_begin
    arm, cold, why = None, None, ""
    want = args.workload if args.workload in ("c2", "c1", "small", "c5") else "c2"
    try:
        if forced:
            arm, why = CpuArm(forced, kind), f"EREF_BENCH_CPU_SAMPLE={forced}"
            cold = arm.run_once()                          # index build, untimed setup
            arm.run_once()                                 # first warm-up run (an extra untimed one when --warmup 0)
        elif want in ("c1", "small"):
            arm, why = CpuArm(want, kind), "asked for (--workload)"
            cold = arm.run_once()
            arm.run_once()
        else:
            # configs[0] first: its times give the estimate, and with its index on disk it is the fallback
            fallback = CpuArm("c1", kind)
            arm = fallback
            c1_cold = fallback.run_once()
            c1_warm = fallback.run_once()
            cold = c1_cold
            reserve = (runs - 1) * c1_warm * 1.1 + 15      # what the fallback still needs from any later moment
            if (kind or ("reference" if have_reference() else "port")) == "reference" and \
                    (_disk_free_gb(tempfile.gettempdir()) < 12 or _host_mem_gb() < 40):
                why = "the headline workload needs 12 GB of scratch disk and 40 GB of RAM for the reference"
            else:
                head = None
                try:
                    head = CpuArm(want, kind)
                    allow = limit - elapsed() - reserve
                    if allow < 30:
                        raise subprocess.TimeoutExpired("eref", allow)
                    head_cold = head.run_once(timeout=allow)                 # index build, untimed setup
                    first = head.run_once(timeout=max(limit - elapsed() - reserve, 1))
                    projected = elapsed() + (runs - 1) * first
                    if projected > limit:
                        why = (f"headline workload: index built in {head_cold:.0f} s, {first:.1f} s per process -> {projected:.0f} s "
                               f"for the whole arm, over the limit of {limit:.0f} s; configs[0] instead")
                        head.close()
                    else:
                        why = (f"the headline workload: index built in {head_cold:.0f} s, {first:.1f} s per process -> {projected:.0f} s "
                               f"for the whole arm, within the limit of {limit:.0f} s")
                        arm, cold = head, head_cold
                        fallback.close()
                except subprocess.TimeoutExpired:
                    why = (f"headline workload: its setup (index build + first process) was abandoned {elapsed():.0f} s into the "
                           f"arm — going on would have left less than the {reserve:.0f} s the configs[0] fallback needs within "
                           f"the limit of {limit:.0f} s; configs[0] instead")
                    head.close()
                except Exception as e:                      # (no scratch space, a reference crash ...): the fallback stands
                    why = f"headline workload failed ({str(e)[-200:]}); configs[0] instead"
                    if head is not None:
                        head.close()
        for _ in range(args.warmup - 1):
            arm.run_once()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            arm.run_once()
        dt = time.perf_counter() - t0
        ms = dt / args.steps * 1e3
        value = arm.bases / 1e9 / (dt / args.steps)
        out = {
            "impl": "reference", "metric": METRIC, "value": round(value, 6), "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(ms, 2), "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "u32", "data": "synthetic",
            "config": config_for(WORKLOAD_DESC[arm.sample_name], args.gpus),
            "cpu_baseline": {"value": round(value, 6), "unit": UNIT, "cores": arm.cores, "kind": arm.kind,
                             "sample": arm.sample + "; each step = one whole eref process, index already on disk",
                             "sample_db_bases": arm.bases, "cold_s": round(cold, 2), "chosen_because": why,
                             "arm_wall_s": round(elapsed(), 1),
                             "selected_refs": arm.stdout.count(b"\n") if arm.stdout is not None else None},
            "e2e": {"value": round(value, 6), "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0,
        }
        _emit(json.dumps(out))
    finally:
        if arm is not None:
            arm.close()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="c2", choices=sorted(WORKLOADS))
    ap.add_argument("--strong", default="c3,c4", help="strong-scaling sub-records: comma list of c3, c4, or none")
    ap.add_argument("--reads", default="shared", choices=["shared", "replicated"],
                    help="N>1, device-resident steps: every GPU packs 1/N of the reads (needs peer pointers) or all of them")
    ap.add_argument("--table", default="auto", choices=["auto", "peer", "allgather", "replicated"],
                    help="N>1: how the query table is built (auto: peer stores over CUDA IPC, else NCCL all-gather)")
    ap.add_argument("--no-prune", action="store_true", help="probe all three streams at every position")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-parity", action="store_true")
    ap.add_argument("--no-unpruned", action="store_true")
    ap.add_argument("--no-c5", action="store_true")
    ap.add_argument("--no-exe-check", action="store_true")
    ap.add_argument("--chunks", type=int, default=4, help="database pieces of the pipelined host-buffer call (e2e)")
    args = ap.parse_args()
    if args.warmup < 3 and args.impl == "b200":
        print(f"bench.py: warmup {args.warmup} < 3 breaks the timing rules", file=sys.stderr)
    # stdout carries the ONE JSON line and nothing else: whatever libraries print there (NCCL's version banner,
    # the reference's own result lines) goes to stderr while the run is in progress
    sys.stdout.flush()
    json_fd = os.dup(1)
    os.dup2(2, 1)
    global _emit
    _emit = lambda line: os.write(json_fd, (line + "\n").encode())
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
