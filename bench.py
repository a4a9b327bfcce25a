#!/usr/bin/env python3
"""bench.py -- BC7 encode throughput of the B200 path, next to the reference's CPU encoder.

Workload (BASELINE.json configs[2], the one `metric` is quoted on): BC7 RGBA encode of a 4096x4096
synthetic albedo with its full mip chain (11 levels, 22,369,616 px, 1,398,101 blocks) at the reference's
quality preset (bc7enc uber level 4, perceptual).  One "step" = one pass over the whole chain.

  python bench.py [--gpus N] [--steps K] [--warmup W]          this repo's CUDA path (through the C ABI)
  python bench.py --impl reference ...                          the reference's own bc7enc on the host cores

value  : Mpix/s, inputs resident in HBM, CUDA-event time of the launches, max over ranks
e2e    : same metric with HOST (pinned) buffers, every step's H2D + kernel + D2H in the timed region, steps back to back through
         hkbcn_import_batch (two pipeline slots); e2e_single_call: one blocking hkbcn_compress_chain() per step
N > 1  : launched by torchrun, one rank per GPU, every rank encodes its own texture (weak scaling, no collective
         on the data path; torch.distributed is used for the barrier and the max-over-ranks only)
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(ROOT, "hork-source_b200", "python"))
sys.path.insert(0, os.path.join(ROOT, "tests"))

BC7 = 66
SIZE = 4096
SEED = 0xB2000004
METRIC = "bc7_encode_mpix_per_s"


# ----------------------------------------------------------------------------- workload
def make_level0(seed, size=SIZE):
    """level 0 = synth.albedo_rgba8 (SURVEY.md 8d 'albedo-B').  Cached in /tmp between the two arms of one round-end run."""
    import synth
    cache = f"/tmp/hkbcn_bench_{seed:x}_{size}.npy"
    if os.path.exists(cache):
        return np.load(cache)
    lvl0 = synth.albedo_rgba8(size, size, seed).reshape(size, size, 4)
    try:
        np.save(cache + ".tmp.npy", lvl0)
        os.replace(cache + ".tmp.npy", cache)
    except OSError:
        pass
    return lvl0


def split_levels(chain_u8, size=SIZE):
    levels, off, n = [], 0, size
    while True:
        levels.append(chain_u8[off:off + n * n * 4].reshape(n, n, 4))
        off += n * n * 4
        if n == 4:
            return levels
        n //= 2


def make_chain_gpu(seed, size=SIZE):
    """the importers' chain (Image.cpp:1755-1803: level i from level i-1, Mitchell / wrap, alpha-weighted, linear) built by
    this library's device-side cascade -- bit-identical to stbir for linear formats (tests/test_gpu_mips.py)"""
    import hkbcn
    lvl0 = make_level0(seed, size)
    mips = hkbcn.lib().hkbcn_calc_num_mips(size, size)
    return split_levels(hkbcn.generate_mips(BC7, lvl0, size, size, mips), size)


def make_chain_ref(seed, size=SIZE):
    """the same chain built by the reference's own stbir_resize (reference arm only)"""
    import refshim
    levels = [make_level0(seed, size)]
    while levels[-1].shape[0] > 4:
        n = levels[-1].shape[0] // 2
        levels.append(refshim.stbir_resize(levels[-1], n, n, alpha_channel=3, flags=0, edge=3, filt=5, space=0))
    return levels


def chain_pixels(levels):
    return int(sum(l.shape[0] * l.shape[1] for l in levels))


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md recipe)."""
    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.rows, self.proc, self.index = [], None, index
        self.t0 = self.t1 = self.tw = None

    def warm(self):
        self.tw = time.perf_counter()

    def begin(self):
        self.t0 = time.perf_counter()

    def end(self):
        self.t1 = time.perf_counter()

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "20"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.perf_counter(), [x.strip() for x in line.split(",")]))

    def stop(self):
        if self.proc:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=5)
            except subprocess.TimeoutExpired:
                self.proc.kill()
        sm, mx, reasons = [], 0, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        # nvidia-smi needs 0.1 - 0.5 s to come up (longer on an 8-GPU box), so it is started before the warm-up and the rows
        # are cut to the timed region by arrival time; a region too short for two rows falls back to warm-up + timed rows
        # (the same kernel under the same load) and says so
        rows = [r for t, r in self.rows if self.t0 is None or (self.t0 <= t <= (self.t1 or t))]
        window = "timed region"
        if len(rows) < 2:
            rows, window = [r for t, r in self.rows if self.tw is None or (self.tw <= t <= (self.t1 or t))], "warm-up + timed region"
        for r in rows:
            try:
                sm.append(float(r[0]))
                mx = max(mx, float(r[1]))
            except (ValueError, IndexError):
                continue
            for n, v in zip(names, r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx or None, "reasons": sorted(reasons), "samples": len(sm),
                "window": window}


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    except (OSError, KeyError, ValueError):
        return 6650.0, "fallback"


# ----------------------------------------------------------------------------- reference arm (CPU)
REF_BUDGET_S = 200.0  # the whole `--impl reference --steps K --warmup W` run should end within a few minutes


def time_reference(levels, steps, warmup, keep_blocks=False):
    """the reference's own bc7enc_compress_block (oracle/_ref, unmodified sources) over all host threads,
    contiguous block ranges per thread (the decomposition CompressBC6h uses, ImageEncoders.cpp:426-486).
    A step is the WHOLE chain (level 0 included) whenever steps + warmup of it fit REF_BUDGET_S on this host -- probed on
    mip levels 1.. (25 % of the pixels) -- else that 25 % sample.  Returns (Mpix/s, s per step, threads, sample text, blocks)."""
    import refshim
    if not refshim.have_ref():
        raise SystemExit("oracle/_ref/libhkref.so missing: run `python __graft_entry__.py build` where /root/reference exists")
    cores = refshim.hw_threads()

    def one(sample):
        return [refshim.compress(BC7, l, l.shape[1], l.shape[0], cores) for l in sample]
    t0 = time.perf_counter()
    one(levels[1:])  # probe (doubles as a warm-up: page faults, thread start-up)
    probe = time.perf_counter() - t0
    whole = probe * chain_pixels(levels) / chain_pixels(levels[1:]) * (steps + warmup) <= REF_BUDGET_S
    sample = levels if whole else levels[1:]
    px = chain_pixels(sample)
    for _ in range(warmup):
        one(sample)
    blocks = None
    t0 = time.perf_counter()
    for _ in range(steps):
        blocks = one(sample)
    dt = (time.perf_counter() - t0) / steps
    text = (f"the whole chain, levels 0..{len(levels) - 1} ({px} px)" if whole else
            f"mip levels 1..{len(levels) - 1} of the chain ({px} px of {chain_pixels(levels)}; the whole chain would not fit {REF_BUDGET_S:.0f} s on {cores} threads)")
    return px / dt / 1e6, dt, cores, text, (blocks if keep_blocks and whole else None)


def cpu_baseline_dict(v, cores, sample):
    return {"value": v, "unit": "Mpix/s", "cores": cores, "per_thread": v / cores, "kind": "reference", "sample": sample}


def run_reference(args):
    if int(os.environ.get("RANK", "0")) != 0:
        return
    levels = make_chain_ref(SEED)
    v, dt, cores, sample, _ = time_reference(levels, max(1, args.steps), args.warmup)
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": "Mpix/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "i32+f32", "data": "synthetic",
        "config": workload_config(levels, args.gpus),
        "cpu_baseline": cpu_baseline_dict(v, cores, sample),
        "e2e": {"value": v, "unit": "Mpix/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }))


def workload_config(levels, n):
    return {"workload": f"BC7_UNORM encode, {SIZE}x{SIZE} RGBA8 synthetic albedo + full mip chain ({len(levels)} levels, {chain_pixels(levels)} px), "
                        "bc7enc uber level 4 perceptual preset (Hk::TextureBlockCompression::CompressBC7)",
            "textures_per_gpu": 1, "parallelism": f"{n} independent textures, one per GPU, no collective",
            "mips": "the importers' cascade: stbir Mitchell / wrap, alpha-weighted, linear (ImageMipmapConfig defaults)", "l2": "256 MiB buffer written between timed steps"}


# ----------------------------------------------------------------------------- sharding (no data-path collective)
def rank_seed(rank):
    """textures are independent units: rank r encodes its own texture, generated from seed SEED + r"""
    return SEED + rank


class Ranks:
    """torch.distributed plumbing only: barrier and max-over-ranks of the timed region.  nccl on GPUs, gloo on CPU
    (tests/test_bench_shard_cpu.py runs the same code path with world_size 2 on CPU)."""

    def __init__(self, use_cuda):
        self.rank = int(os.environ.get("RANK", "0"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.use_cuda = use_cuda
        self.dist = None
        if self.world > 1:
            import torch
            import torch.distributed as dist
            if use_cuda:
                dist.init_process_group("nccl", device_id=torch.device("cuda", self.local))
                # host-side waits (rank 0 driving all GPUs through the library while the others stand by) must not put a
                # spinning NCCL kernel on the other ranks' GPUs: a gloo group carries those
                self.host_group = dist.new_group(backend="gloo")
            else:
                dist.init_process_group("gloo")
                self.host_group = None
            self.dist = dist

    def barrier(self):
        if self.dist:
            self.dist.barrier()
        if self.use_cuda:
            import torch
            torch.cuda.synchronize()

    def host_barrier(self):
        """a barrier that leaves the GPUs alone"""
        if self.dist:
            self.dist.barrier(group=self.host_group) if self.host_group is not None else self.dist.barrier()

    def max(self, x):
        import torch
        t = torch.tensor([x], dtype=torch.float64, device="cuda" if self.use_cuda else "cpu")
        if self.dist:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def sum(self, x):
        import torch
        t = torch.tensor([x], dtype=torch.float64, device="cuda" if self.use_cuda else "cpu")
        if self.dist:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.SUM)
        return float(t.item())

    def close(self):
        if self.dist:
            self.dist.destroy_process_group()


def whole_job_value(ranks, px_this_rank, ms_this_rank):
    """whole-job Mpix/s = pixels all ranks processed / max-over-ranks time"""
    ms = ranks.max(ms_this_rank)
    px = ranks.sum(px_this_rank)
    return px / (ms * 1e-3) / 1e6, ms


def run_selftest(args):
    """CPU check of the N > 1 plumbing (gloo): every rank derives its shard, reports a synthetic time, rank 0 prints
    the aggregate exactly as the CUDA arm does.  No kernels, no oracle."""
    ranks = Ranks(use_cuda=False)
    px = 1000 * (ranks.rank + 1)
    ms = 10.0 * (ranks.rank + 1)
    ranks.barrier()
    value, ms_max = whole_job_value(ranks, px, ms)
    if ranks.rank == 0:
        print(json.dumps({"selftest": True, "n_gpus": ranks.world, "value": value, "ms_per_step": ms_max,
                          "seeds": [rank_seed(r) for r in range(ranks.world)]}))
    ranks.close()


# ----------------------------------------------------------------------------- CUDA arm
def run_gpu(args):
    import torch
    import hkbcn

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- this path has no CPU fallback")
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
    ranks = Ranks(use_cuda=True)
    rank, local, world, dist = ranks.rank, ranks.local, ranks.world, ranks.dist
    barrier, max_over_ranks = ranks.barrier, ranks.max

    L = hkbcn.lib()
    hkbcn.init(1 << local)  # library device 0 == cuda:local
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()  # long before the timed region: nvidia-smi takes a while to come up
    levels = make_chain_gpu(rank_seed(rank))
    mips, px = len(levels), chain_pixels(levels)
    chain_identity = None
    if rank == 0:
        # the reference arm encodes the chain the reference's own stbir builds; this arm's chain comes from the device cascade.
        # They must be the same bytes or the two arms do not run the same workload.
        import refshim
        if refshim.have_ref():
            ref_levels = make_chain_ref(rank_seed(rank))
            chain_identity = all(np.array_equal(a, b) for a, b in zip(levels, ref_levels))
            if not chain_identity:
                raise SystemExit("bench.py: the device mip cascade's chain differs from the reference's stbir chain")
    src_np = np.concatenate([l.ravel() for l in levels])
    dst_bytes = L.hkbcn_chain_bytes(BC7, SIZE, SIZE, 1, mips)
    assert dst_bytes == px  # 16 B per 16 px

    # device-resident input / output (torch = allocator + stream/event plumbing only)
    d_src = torch.from_numpy(src_np).cuda()
    d_dst = torch.empty(dst_bytes, dtype=torch.uint8, device="cuda")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    stream = torch.cuda.current_stream()

    def step():
        hkbcn.check(L.hkbcn_compress_chain_dev(0, BC7, d_src.data_ptr(), d_dst.data_ptr(), SIZE, SIZE, mips, 0, C.c_void_p(stream.cuda_stream)))

    sampler.warm()
    for i in range(max(args.warmup, 3)):
        step()
    barrier()
    launches0 = L.hkbcn_launch_count()
    evs = []
    barrier()
    sampler.begin()
    for i in range(args.steps):
        flush.fill_(i & 255)  # evict the input from the 126 MB L2 (untimed)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        step()
        e1.record(stream)
        evs.append((e0, e1))
    barrier()
    sampler.end()
    launches = L.hkbcn_launch_count() - launches0
    value, ms_step = whole_job_value(ranks, px, sum(a.elapsed_time(b) for a, b in evs) / args.steps)
    clocks = sampler.stop() if rank == 0 else None

    # ---- end to end through the C ABI with host buffers (H2D + kernel + D2H inside the timed region)
    h_src = L.hkbcn_alloc_pinned(src_np.nbytes)
    h_dst = L.hkbcn_alloc_pinned(dst_bytes)
    if not h_src or not h_dst:
        raise SystemExit("pinned allocation failed: " + L.hkbcn_last_error().decode())
    C.memmove(h_src, src_np.ctypes.data, src_np.nbytes)
    e2e_steps = max(1, min(args.steps, 10))

    def e2e_step():
        hkbcn.check(L.hkbcn_compress_chain(BC7, h_src, h_dst, SIZE, SIZE, mips, 0))
    for _ in range(2):
        e2e_step()
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        e2e_step()
    torch.cuda.synchronize()
    e2e_s = max_over_ranks(time.perf_counter() - t0) / e2e_steps
    barrier()
    # the two paths must agree byte for byte
    host_out = np.ctypeslib.as_array(C.cast(h_dst, C.POINTER(C.c_uint8)), shape=(dst_bytes,))
    if not np.array_equal(host_out, d_dst.cpu().numpy()):
        raise SystemExit("bench.py: host-pointer and device-pointer encodes differ")
    e2e_value = world * px / e2e_s / 1e6

    # ---- the same steps back to back through the batch entry point: hkbcn_import_batch over K jobs that each hand over the whole
    # pre-built chain (HKBCN_FLAG_SRC_CHAIN, the same 89.5 MB per step as above, pinned) and receive the compressed chain in
    # memory.  The library's two pipeline slots put the upload of step i + 1 and the download of step i - 1 under the kernel
    # of step i; every step's copies are inside the timed region.
    pipe_k = e2e_steps
    pipe_dst = [L.hkbcn_alloc_pinned(dst_bytes) for _ in range(3)]  # two slots in flight: three buffers never collide
    if not all(pipe_dst):
        raise SystemExit("pinned allocation failed: " + L.hkbcn_last_error().decode())
    pjobs = (hkbcn.ImportJob * pipe_k)()
    for i in range(pipe_k):
        pjobs[i].fmt, pjobs[i].width, pjobs[i].height, pjobs[i].generate_mips = BC7, SIZE, SIZE, 0
        pjobs[i].flags = 1 << 18  # HKBCN_FLAG_SRC_CHAIN
        pjobs[i].src_level0, pjobs[i].out_path, pjobs[i].dst = h_src, None, pipe_dst[i % 3]
    hkbcn.check(L.hkbcn_import_batch(pjobs, min(pipe_k, 3)))  # warm-up: slot buffers, streams
    barrier()
    t0 = time.perf_counter()
    hkbcn.check(L.hkbcn_import_batch(pjobs, pipe_k))
    pipe_s = max_over_ranks(time.perf_counter() - t0) / pipe_k
    barrier()
    for b in pipe_dst[: min(pipe_k, 3)]:
        if not np.array_equal(np.ctypeslib.as_array(C.cast(b, C.POINTER(C.c_uint8)), shape=(dst_bytes,)), host_out):
            raise SystemExit("bench.py: pipelined and single-call encodes differ")
    pipe_value = world * px / pipe_s / 1e6

    # ---- the same call on PAGEABLE host memory: what the drop-in caller hands over (Image.cpp passes HeapBlob memory)
    pg_dst = np.zeros(dst_bytes, np.uint8)

    def pageable_step():
        hkbcn.check(L.hkbcn_compress_chain(BC7, src_np.ctypes.data_as(C.c_void_p), pg_dst.ctypes.data_as(C.c_void_p), SIZE, SIZE, mips, 0))
    for _ in range(2):
        pageable_step()
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        pageable_step()
    torch.cuda.synchronize()
    pg_s = max_over_ranks(time.perf_counter() - t0) / e2e_steps
    barrier()
    if not np.array_equal(pg_dst, host_out):
        raise SystemExit("bench.py: pageable and pinned host-pointer encodes differ")
    pg_value = world * px / pg_s / 1e6

    # ---- the importers' whole cascade through the C ABI: level 0 in (host), mips generated on the device, chain encoded,
    # compressed chain out (host).  Only level 0 crosses PCIe.  Same bytes out as the pre-built chain (the cascade is
    # bit-identical to stbir for linear formats).
    l0_bytes = SIZE * SIZE * 4

    def import_step():
        hkbcn.check(L.hkbcn_import_chain(BC7, h_src, h_dst, SIZE, SIZE, mips, None, 0))
    for _ in range(2):
        import_step()
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        import_step()
    torch.cuda.synchronize()
    imp_s = max_over_ranks(time.perf_counter() - t0) / e2e_steps
    barrier()
    if not np.array_equal(host_out, d_dst.cpu().numpy()):
        raise SystemExit("bench.py: import cascade and pre-built chain encodes differ")
    import_value = world * px / imp_s / 1e6

    # ---- BASELINE.json configs[4]: a batch of mixed BC1 / BC5 / BC7 2048^2 textures with full mip chains through
    # hkbcn_import_batch (level 0 in pinned host memory -> upload, device mip cascade, encode, download; two pipeline slots
    # per GPU).  Every rank imports its own batch on its own GPU; no collective.
    batch = batch_import(hkbcn, L, rank)
    barrier()
    batch_s = max_over_ranks(batch["s"])
    batch_value = world * batch["mpix"] / batch_s

    if rank == 0:
        hbm, which = peaks()
        hbm_achieved = 5.0 * px / (ms_step * 1e-3) / 1e9  # algorithmic bytes: 4 B/px read + 1 B/px written
        out = {
            "metric": METRIC, "value": value, "unit": "Mpix/s", "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "i32+f32", "data": "synthetic",
            "config": workload_config(levels, world),
            "e2e": {"value": pipe_value, "unit": "Mpix/s", "h2d_bytes_per_step": int(src_np.nbytes), "d2h_bytes_per_step": int(dst_bytes),
                    "ms_per_step": pipe_s * 1e3, "steps": pipe_k,
                    "api": "hkbcn_import_batch, one job per step (HKBCN_FLAG_SRC_CHAIN: the pre-built chain in pinned host memory in, the "
                           "compressed chain in pinned host memory out), steps back to back: the library's two pipeline slots overlap a "
                           "step's copies with its neighbours' kernels"},
            "e2e_single_call": {"value": e2e_value, "unit": "Mpix/s", "h2d_bytes_per_step": int(src_np.nbytes), "d2h_bytes_per_step": int(dst_bytes),
                                "ms_per_step": e2e_s * 1e3, "api": "hkbcn_compress_chain (host pointers, pinned), one blocking call per step: "
                                "upload, kernel, download in sequence (round 1's e2e)"},
            "e2e_pageable": {"value": pg_value, "unit": "Mpix/s", "ms_per_step": pg_s * 1e3,
                             "api": "hkbcn_compress_chain on malloc'd (pageable) buffers -- what Image.cpp's HeapBlob memory is to the runtime"},
            "import_chain": {"value": import_value, "unit": "Mpix/s", "ms_per_step": imp_s * 1e3, "h2d_bytes_per_step": l0_bytes,
                             "d2h_bytes_per_step": int(dst_bytes), "api": "hkbcn_import_chain (level 0 in, device mip cascade + encode, chain out)"},
            "batch": {"value": batch_value, "unit": "Mpix/s", "textures_per_gpu": batch["n"], "mix": batch["mix"], "s": batch_s,
                      "h2d_bytes": batch["h2d"], "d2h_bytes": batch["d2h"],
                      "api": "hkbcn_import_batch (pinned level 0 in, compressed chains out, mips on the device)"},
            "gpu_launches": int(launches), "clocks": clocks,
            "roofline": roofline_block(px, ms_step, clocks, hbm, which, hbm_achieved),
            "chain_equals_stbir_chain": chain_identity,
        }
        if world == 1:
            import refshim
            out["decode"] = decode_sweep(torch, hkbcn, L, d_dst, stream, flush, hbm)
            out["encode_other_formats"] = encode_others(torch, hkbcn, L, stream, flush)
            out["next_rows"] = next_rows(torch, hkbcn, L, stream, flush, hbm)
            out["cube_bc6h"] = cube_bc6h(torch, hkbcn, L, stream, flush)
            # (ii) of BASELINE.md 3: the reference's per-block encoder over all host threads, on the whole chain, once; its
            # blocks are kept and compared with the GPU's
            v, dt, cores, sample, ref_blocks = time_reference(levels, 1, 0, keep_blocks=True)
            out["cpu_baseline"] = cpu_baseline_dict(v, cores, sample)
            out["parity"] = parity_block(levels, host_out, ref_blocks)
            # (i) as shipped: CompressBC7 is a single-threaded loop (ImageEncoders.cpp:524-557); bounded: the 1024^2 level
            lv = levels[2]
            t0 = time.perf_counter()
            refshim.compress(BC7, lv, lv.shape[1], lv.shape[0], 1)
            dt1 = time.perf_counter() - t0
            out["cpu_as_shipped"] = {"value": lv.shape[0] * lv.shape[1] / dt1 / 1e6, "unit": "Mpix/s", "cores": 1, "kind": "reference",
                                     "sample": f"CompressBC7 (one thread, as the reference ships it) on the {lv.shape[1]}^2 level"}
            # the CPU side of the import cascade: the reference's stbir_resize is single-threaded as shipped
            t0 = time.perf_counter()
            cur = levels[0]
            while cur.shape[0] > 4:
                cur = refshim.stbir_resize(cur, cur.shape[1] // 2, cur.shape[0] // 2, alpha_channel=3, flags=0, edge=3, filt=5, space=0)
            out["import_chain"]["cpu_stbir_cascade_ms"] = (time.perf_counter() - t0) * 1e3
    L.hkbcn_free_pinned(h_src)
    L.hkbcn_free_pinned(h_dst)
    # ---- the library's own sharding (BASELINE.json configs[4]): rank 0 alone re-initialises the library over
    # all N GPUs of the job and drives them through hkbcn_import_batch / hkbcn_compress_chain_multi; the other ranks stand by
    del d_src, d_dst, flush
    torch.cuda.empty_cache()
    ranks.host_barrier()
    if rank == 0:
        try:
            out["batch_sharded"], out["banded"] = library_sharding(torch, hkbcn, L, world, levels)
        except Exception as ex:  # the headline line must survive a failure of this extra
            out["batch_sharded"] = {"error": repr(ex)}
        print(json.dumps(out))
    ranks.host_barrier()
    ranks.close()


def device_texture(torch, dev, kind, n, seed):
    """a distinct synthetic n x n source generated ON the GPU (no PCIe): smooth value noise from a random lattice upsampled
    bilinearly over 4 octaves, hard-edged cells, fine grain -- the ingredients of python/synth.py, from torch's device RNG.
    kind: 'albedo' RGBA8 (alpha 255), 'normal' RG8 (tangent-space normals of a noise height field)"""
    g = torch.Generator(device=dev).manual_seed(seed)
    F = torch.nn.functional

    def noise(ch, octaves=4, base=32):
        acc, amp, tot = 0, 1.0, 0.0
        for o in range(octaves):
            cells = max(n // (base >> o), 2)
            lat = torch.rand((1, ch, cells, cells), device=dev, generator=g)
            acc = acc + amp * F.interpolate(lat, size=(n, n), mode="bilinear", align_corners=False)
            tot += amp
            amp *= 0.5
        return (acc / tot)[0]
    if kind == "albedo":
        luma = noise(1)
        img = (0.65 * luma + 0.35 * noise(3) + 0.04 * (torch.rand((3, n, n), device=dev, generator=g) - 0.5)).clamp(0, 1)
        cell = torch.rand((1, 4, n // 16, n // 16), device=dev, generator=g)
        cell = F.interpolate(cell, size=(n, n), mode="nearest")[0]
        flat = cell[3:4] < 0.05
        img = torch.where(flat, cell[:3], img)
        rgba = torch.cat([img, torch.ones((1, n, n), device=dev)], 0)
        return (rgba * 255.0 + 0.5).to(torch.uint8).permute(1, 2, 0).contiguous()
    h = noise(1, 5, 64)[0] * 24.0
    dx = (torch.roll(h, -1, 1) - torch.roll(h, 1, 1)) * 0.5
    dy = (torch.roll(h, -1, 0) - torch.roll(h, 1, 0)) * 0.5
    inv = torch.rsqrt(dx * dx + dy * dy + 1.0)
    rg = torch.stack([(-dx * inv) * 0.5 + 0.5, (-dy * inv) * 0.5 + 0.5], -1).clamp(0, 1)
    return (rg * 255.0 + 0.5).to(torch.uint8).contiguous()


def library_sharding(torch, hkbcn, L, world, levels, per_gpu=256, n=2048):
    """rank 0, library initialised over `world` GPUs:
    batch_sharded -- per_gpu x world textures (50 % BC1_UNORM_SRGB, 25 % BC5_UNORM, 25 % BC7_UNORM, 2048^2 with full chains),
                     every source distinct and generated on the GPU the library's plan assigns it to (HKBCN_FLAG_SRC_DEVICE),
                     compressed chains into pinned host memory; whole textures per GPU, no collective
    banded        -- ONE 4096^2 BC7 chain over the GPUs in bands of block rows (hkbcn_compress_chain_multi)"""
    L.hkbcn_shutdown()
    hkbcn.init((1 << world) - 1)
    assert L.hkbcn_device_count() == world
    mips = L.hkbcn_calc_num_mips(n, n)
    total = per_gpu * world
    fmts = [(55, "albedo", 5), (62, "normal", 2), (55, "albedo", 5), (66, "albedo", 5)]  # i % 4 -> 50 / 25 / 25
    arr = (hkbcn.ImportJob * total)()
    for i in range(total):
        fmt, kind, filt = fmts[i % 4]
        arr[i] = hkbcn.ImportJob(fmt, n, n, 1, hkbcn.MipConfig(3, filt, 0, 0), hkbcn.FLAG_SRC_DEVICE, None, None, None, 0)
    plan = (C.c_int * total)()
    hkbcn.check(L.hkbcn_plan_import_batch(arr, C.c_size_t(total), world, plan))
    srcs, pinned, px = [], [], 0
    t0 = time.perf_counter()
    for i in range(total):
        fmt, kind, filt = fmts[i % 4]
        t = device_texture(torch, torch.device("cuda", plan[i]), kind, n, 0xB2100000 + i)
        srcs.append(t)
        nb = L.hkbcn_chain_bytes(fmt, n, n, 1, mips)
        h_dst = L.hkbcn_alloc_pinned(nb)
        if not h_dst:
            raise RuntimeError("pinned allocation failed")
        pinned.append(h_dst)
        arr[i].src_level0 = t.data_ptr()
        arr[i].dst = h_dst
        px += sum(max(n >> l, 4) ** 2 for l in range(mips))
    for d in range(world):
        torch.cuda.synchronize(d)
    gen_s = time.perf_counter() - t0
    best, stats = None, None
    for it in range(2):  # the first pass warms buffers and tap tables
        t0 = time.perf_counter()
        hkbcn.check(L.hkbcn_import_batch(arr, C.c_size_t(total)))
        dt = time.perf_counter() - t0
        if it:
            best = dt
            busy = (C.c_double * world)()
            L.hkbcn_batch_stats(busy, world)
            stats = [round(b, 1) for b in busy]
    for i in range(total):
        if arr[i].status:
            raise RuntimeError(f"import job {i} failed: {arr[i].status}")
    for p in pinned:
        L.hkbcn_free_pinned(p)
    del srcs
    per_dev = [sum(1 for i in range(total) if plan[i] == d) for d in range(world)]
    sharded = {"value": px / best / 1e6, "unit": "Mpix/s", "textures": total, "per_gpu": per_dev, "s": best, "busy_ms_per_gpu": stats,
               "lpt_imbalance": round(max(stats) / (sum(stats) / len(stats)), 3), "source": "generated on the assigned GPU (torch device RNG), distinct per texture",
               "generate_s": round(gen_s, 2), "mix": "50 % BC1_UNORM_SRGB / 25 % BC5_UNORM / 25 % BC7_UNORM, 2048^2 + full chains",
               "d2h_bytes": int(sum(L.hkbcn_chain_bytes(fmts[i % 4][0], n, n, 1, mips) for i in range(total))), "h2d_bytes": 0,
               "api": "hkbcn_plan_import_batch + hkbcn_import_batch (HKBCN_FLAG_SRC_DEVICE), one process driving all GPUs"}
    # one 4096^2 BC7 chain in bands over the GPUs (pinned host buffers), byte-compared with the single-GPU result
    src_np = np.concatenate([l.ravel() for l in levels])
    nbytes = L.hkbcn_chain_bytes(BC7, SIZE, SIZE, 1, len(levels))
    h_src, h_dst = L.hkbcn_alloc_pinned(src_np.nbytes), L.hkbcn_alloc_pinned(nbytes)
    C.memmove(h_src, src_np.ctypes.data, src_np.nbytes)
    times = {}
    ref_out = None
    for g in sorted({1, world}):
        ts = []
        for it in range(4):
            t0 = time.perf_counter()
            hkbcn.check(L.hkbcn_compress_chain_multi(BC7, h_src, h_dst, SIZE, SIZE, len(levels), 0, g))
            ts.append(time.perf_counter() - t0)
        times[g] = min(ts[1:])
        got = np.ctypeslib.as_array(C.cast(h_dst, C.POINTER(C.c_uint8)), shape=(nbytes,)).copy()
        if ref_out is None:
            ref_out = got
        elif not np.array_equal(got, ref_out):
            raise RuntimeError("banded chain differs from the single-GPU chain")
    L.hkbcn_free_pinned(h_src)
    L.hkbcn_free_pinned(h_dst)
    pxc = chain_pixels(levels)
    banded = {"value": pxc / times[world] / 1e6, "unit": "Mpix/s", "gpus": world, "ms": round(times[world] * 1e3, 2), "ms_one_gpu": round(times[1] * 1e3, 2),
              "identical_to_one_gpu": True, "api": "hkbcn_compress_chain_multi (host pointers, pinned): bands of block rows per GPU, no exchange"}
    return sharded, banded


def batch_import(hkbcn, L, rank, per_format=8, n=2048):
    """per_format x (BC1_UNORM_SRGB, BC5_UNORM, BC7_UNORM) textures of n x n with full chains; returns timing of one batch"""
    import synth
    mips = L.hkbcn_calc_num_mips(n, n)
    srcs = {55: synth.albedo_rgba8(n, n, 0xB2000010 + rank, "opaque"), 62: synth.normals_rg8(n, n, 0xB2000011 + rank),
            66: synth.albedo_rgba8(n, n, 0xB2000012 + rank, "opaque")}
    pinned, jobs = [], []
    arr = (hkbcn.ImportJob * (3 * per_format))()
    h2d = d2h = 0
    for fmt, img in srcs.items():
        img = np.ascontiguousarray(img).view(np.uint8).ravel()
        h_src = L.hkbcn_alloc_pinned(img.nbytes)
        C.memmove(h_src, img.ctypes.data, img.nbytes)
        pinned.append(h_src)
        nb = L.hkbcn_chain_bytes(fmt, n, n, 1, mips)
        cfg = hkbcn.MipConfig(3, 2 if fmt == 62 else 5, 0, 0)  # normal maps: triangle (Image.cpp:2359); albedo: Mitchell
        for _ in range(per_format):
            h_dst = L.hkbcn_alloc_pinned(nb)
            pinned.append(h_dst)
            jobs.append((fmt, cfg, h_src, h_dst))
            h2d += img.nbytes
            d2h += nb
    order = [jobs[i % 3 * per_format + i // 3] for i in range(len(jobs))]  # interleave the formats as a mixed manifest would
    for i, (fmt, cfg, h_src, h_dst) in enumerate(order):
        arr[i] = hkbcn.ImportJob(fmt, n, n, 1, cfg, 0, h_src, None, h_dst, 0)
    best = None
    for it in range(3):  # first pass warms buffers and tap tables
        t0 = time.perf_counter()
        hkbcn.check(L.hkbcn_import_batch(arr, C.c_size_t(len(order))))
        dt = time.perf_counter() - t0
        if it and (best is None or dt < best):
            best = dt
    for p in pinned:
        L.hkbcn_free_pinned(p)
    lv = sum(max(n >> i, 4) ** 2 for i in range(mips))
    return {"s": best, "mpix": len(order) * lv / 1e6, "n": len(order), "mix": f"{per_format} x BC1_UNORM_SRGB + {per_format} x BC5_UNORM + "
            f"{per_format} x BC7_UNORM, {n}x{n} + mips", "h2d": int(h2d), "d2h": int(d2h)}


def _timed(torch, stream, flush, fn, iters, skip=2):
    times = []
    for i in range(iters + skip):
        flush.fill_(i & 255)  # evict L2 between iterations (untimed)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        fn()
        e1.record(stream)
        torch.cuda.synchronize()
        if i >= skip:
            times.append(e0.elapsed_time(e1))
    return float(np.mean(times))


def decode_sweep(torch, hkbcn, L, bc7_blocks, stream, flush, hbm):
    """HBM-bound half of the path (BASELINE.json configs[4] 'decode bandwidth sweep', SURVEY.md 8d: sizes 256^2 .. 8192^2, all
    seven formats): whole-surface decode per format and size (16384^2 kept as the asymptote), algorithmic bytes/px of SURVEY.md
    8d (block bytes read + texel bytes written).  `batched`: 2048^2 surfaces as ONE storage of slices (hkbcn_decompress_slices_dev,
    one launch) sized to >= 64 MB of output -- what a texture set looks like to the path; a lone small surface is a 10 us
    kernel whose launch, ramp and tail no memory system can hide.  BC7 and BC6H inputs are blocks this library's encoders
    produced (the mode mix of real content), tiled up to size; the others are uniform random bits (every selector and both
    palette modes).  `cpu`: the reference's decoders (bcdec / bc7decomp) over all host threads at 4096^2."""
    import refshim
    import synth
    out = {}
    gen = torch.Generator(device="cuda").manual_seed(7)
    base7 = bc7_blocks[: (SIZE // 4) ** 2 * 16].view(SIZE // 4, SIZE // 4, 16)
    base6 = {}
    for fmt in (64, 65):
        img = synth.sky_rgba32f(256, 256, 0xB2000006, 0)
        if fmt == 65:
            img[..., :3] -= 1.0
        hdr = torch.from_numpy(np.ascontiguousarray(img).view(np.uint8).ravel()).cuda()
        b6 = torch.empty(64 * 64 * 16, dtype=torch.uint8, device="cuda")
        hkbcn.check(L.hkbcn_compress_dev(0, fmt, hdr.data_ptr(), b6.data_ptr(), 256, 256, 0, C.c_void_p(stream.cuda_stream)))
        torch.cuda.synchronize()
        base6[fmt] = b6.view(64, 64, 16)

    def blocks_for(fmt, blk, nbx, nby):
        if fmt == 66:
            return base7.repeat((nby + base7.shape[0] - 1) // base7.shape[0], (nbx + base7.shape[1] - 1) // base7.shape[1], 1)[:nby, :nbx].contiguous().view(-1)
        if fmt in (64, 65):
            return base6[fmt].repeat((nby + 63) // 64, (nbx + 63) // 64, 1)[:nby, :nbx].contiguous().view(-1)
        return torch.randint(0, 256, (nbx * nby * blk,), dtype=torch.uint8, device="cuda", generator=gen)

    cs = C.c_void_p(stream.cuda_stream)
    for name, fmt, blk, tex, flags in (("BC1", 54, 8, 4, 0), ("BC2", 56, 16, 4, 0), ("BC3", 58, 16, 4, 0), ("BC4", 60, 8, 1, 0), ("BC5", 62, 16, 2, 0),
                                       ("BC7", 66, 16, 4, 0), ("BC6H_f16", 64, 16, 6, 0), ("BC6H_S_f16", 65, 16, 6, 0), ("BC6H_f32", 64, 16, 12, 1 << 8)):
        out[name] = {}
        for n in (256, 512, 1024, 2048, 4096, 8192, 16384):
            nb = n // 4
            src = blocks_for(fmt, blk, nb, nb)
            dst = torch.empty(n * n * tex, dtype=torch.uint8, device="cuda")
            ms = _timed(torch, stream, flush, lambda: hkbcn.check(L.hkbcn_decompress_dev(
                0, fmt, src.data_ptr(), dst.data_ptr(), n, n, n * tex, flags, cs)), 5, 3)
            gbs = (nb * nb * blk + n * n * tex) / (ms * 1e-3) / 1e9
            out[name][str(n)] = {"GB/s": round(gbs, 1), "frac": round(gbs / hbm, 4), "Gpix/s": round(n * n / (ms * 1e-3) / 1e9, 2), "us": round(ms * 1e3, 1)}
            if n == 4096 and refshim.have_ref():
                hb = src.cpu().numpy()
                t0 = time.perf_counter()
                refshim.decompress(fmt, hb, n, n, f32out=bool(flags))
                dt = time.perf_counter() - t0
                out[name]["cpu_4096"] = {"GB/s": round((nb * nb * blk + n * n * tex) / dt / 1e9, 2), "Gpix/s": round(n * n / dt / 1e9, 3), "cores": refshim.hw_threads(),
                                         "kind": "reference"}
            del src, dst
        # a set of 2048^2 textures as one storage of slices, decoded by ONE launch: 64 MB and 256 MB of decoded texels.  What a
        # launch moves decides the fraction, not the shape: 64 MB is ~11 us of HBM time next to ~7 us of launch, ramp and tail
        n = 2048
        for mb in (64, 256):
            slices = max(1, -(-(mb << 20) // (n * n * tex)))
            src = blocks_for(fmt, blk, n // 4, (n // 4) * slices)
            dst = torch.empty(n * n * tex * slices, dtype=torch.uint8, device="cuda")
            ms = _timed(torch, stream, flush, lambda: hkbcn.check(L.hkbcn_decompress_slices_dev(
                0, fmt, src.data_ptr(), dst.data_ptr(), n, n, slices, 1, flags, cs)), 5, 3)
            gbs = slices * ((n // 4) ** 2 * blk + n * n * tex) / (ms * 1e-3) / 1e9
            out[name][f"batched_2048_{mb}MB"] = {"slices": slices, "GB/s": round(gbs, 1), "frac": round(gbs / hbm, 4), "us": round(ms * 1e3, 1),
                                                 "api": "hkbcn_decompress_slices_dev, one launch"}
            del src, dst
    return out


def encode_others(torch, hkbcn, L, stream, flush):
    """the other encoders of the path on BASELINE.json's configs (level 0 only), device-resident, Mpix/s -- each with the
    reference's own encoder over all host threads beside it (whole surface; BC6H: a 256^2 crop, the reference needs ~15 s
    per Mpix on 16 threads) and, for the bit-exact encoders, the count of blocks identical to the reference's"""
    import refshim
    import synth
    out = {}
    for name, fmt, n, img in (("BC1_2048", 54, 2048, lambda: synth.albedo_rgba8(2048, 2048, 0xB2000001)),
                              ("BC3_2048", 58, 2048, lambda: synth.albedo_rgba8(2048, 2048, 0xB2000005, "mask")),
                              ("BC4_4096", 60, 4096, lambda: synth.roughness_r8(4096, 4096, 0xB2000002)),
                              ("BC5_4096", 62, 4096, lambda: synth.normals_rg8(4096, 4096, 0xB2000003)),
                              ("BC7_alpha_2048", 66, 2048, lambda: synth.albedo_rgba8(2048, 2048, 0xB2000005, "mask")),
                              ("BC6H_1024", 64, 1024, lambda: synth.sky_rgba32f(1024, 1024, 0xB2000006, 0)),
                              ("BC6H_512", 64, 512, lambda: synth.sky_rgba32f(512, 512, 0xB2000006, 0))):
        src = np.ascontiguousarray(img())
        d_src = torch.from_numpy(src.view(np.uint8).ravel()).cuda()
        bb = hkbcn.block_bytes(fmt)
        d_dst = torch.empty((n // 4) ** 2 * bb, dtype=torch.uint8, device="cuda")
        ms = _timed(torch, stream, flush, lambda: hkbcn.check(L.hkbcn_compress_dev(
            0, fmt, d_src.data_ptr(), d_dst.data_ptr(), n, n, 0, C.c_void_p(stream.cuda_stream))), 3, 1)
        out[name] = {"Mpix/s": round(n * n / ms / 1e3, 1), "ms": round(ms, 3)}
        if refshim.have_ref() and name != "BC6H_512":
            cores = refshim.hw_threads()
            if fmt == 64:
                m = 256
                crop = np.ascontiguousarray(src.reshape(n, n, 4)[256:256 + m, 512:512 + m])
                t0 = time.perf_counter()
                refshim.compress(fmt, crop, m, m, cores)
                dt = time.perf_counter() - t0
                out[name]["cpu_reference"] = {"Mpix/s": round(m * m / dt / 1e6, 4), "cores": cores, "sample": f"{m}^2 crop of the face"}
            else:
                t0 = time.perf_counter()
                want = refshim.compress(fmt, src, n, n, cores)
                dt = time.perf_counter() - t0
                same = (d_dst.cpu().numpy().reshape(-1, bb) == want.reshape(-1, bb)).all(axis=1)
                out[name]["cpu_reference"] = {"Mpix/s": round(n * n / dt / 1e6, 2), "cores": cores, "sample": "the whole surface"}
                out[name]["identical_blocks"] = [int(same.sum()), int(same.size)]
        del d_src, d_dst
    return out


def cube_bc6h(torch, hkbcn, L, stream, flush, size=1024):
    """BASELINE.json configs[3]: a 6 x 1024^2 RGBA32F sky cube -> GGX-prefiltered chain (max_lod 10, roughness L/10, 1024^2 ..
    1^2, the envprobegen.frag arithmetic) -> BC6H_UFLOAT of every level >= 4^2 of every face in ONE launch
    (hkbcn_compress_slices_dev, ImageStorage order).  Device-resident; the reference's own encoder (CompressBC6h, all host
    threads) is timed beside it on one 128^2 face level of the same chain."""
    import time as _t
    import synth
    import refshim
    cs = C.c_void_p(stream.cuda_stream)
    max_lod = int(np.log2(size))
    mips = max_lod - 1  # 1024 .. 4
    cube = torch.from_numpy(np.stack([synth.sky_rgba32f(size, size, 0xB2000006, f) for f in range(6)])).cuda()
    texels = L.hkbcn_envprobe_texels(max_lod)
    chain = torch.empty(texels * 4, dtype=torch.float32, device="cuda")
    ms_pre = _timed(torch, stream, flush, lambda: hkbcn.check(L.hkbcn_envprobe_prefilter_dev(0, cube.data_ptr(), size, max_lod, chain.data_ptr(), cs)), 2, 1)
    px = sum(6 * (size >> i) ** 2 for i in range(mips))
    dst = torch.empty(L.hkbcn_chain_bytes(64, size, size, 6, mips), dtype=torch.uint8, device="cuda")
    ms_enc = _timed(torch, stream, flush, lambda: hkbcn.check(L.hkbcn_compress_slices_dev(0, 64, chain.data_ptr(), dst.data_ptr(), size, size, 6, mips, 0, cs)), 2, 1)
    out = {"workload": f"6 x {size}^2 RGBA32F sky cube, GGX-prefiltered chain of {mips} levels ({px} px), BC6H_UFLOAT, one launch",
           "prefilter_ms": round(ms_pre, 2), "prefilter_Gsamples/s": round(texels * 1024 / ms_pre / 1e6, 1),
           "encode_ms": round(ms_enc, 1), "encode_Mpix/s": round(px / ms_enc / 1e3, 2),
           "api": "hkbcn_envprobe_prefilter_dev + hkbcn_compress_slices_dev"}
    if refshim.have_ref():
        # bounded CPU sample: face 0 of the 128^2 level of the same prefiltered chain (16384 px), reference CompressBC6h
        lvl = max_lod - 7
        off = sum(6 * (size >> i) ** 2 for i in range(lvl)) * 4
        n = size >> lvl
        face = chain[off:off + n * n * 4].cpu().numpy()
        t0 = _t.perf_counter()
        refshim.compress(64, face, n, n, refshim.hw_threads())
        dt = _t.perf_counter() - t0
        out["cpu_reference"] = {"Mpix/s": round(n * n / dt / 1e6, 5), "cores": refshim.hw_threads(), "kind": "reference",
                                "sample": f"face 0 of the {n}^2 level ({n * n} px), CompressBC6h on all host threads"}
    return out


def next_rows(torch, hkbcn, L, stream, flush, hbm):
    """SURVEY.md 8f rows f4 and f3, device-resident: the normal-map packers / roughness bake (HBM-bound streaming kernels,
    algorithmic bytes = Float3 in + packed texels out) and the environment-probe kernels (gather-bound; samples/s), with the
    C restatement of the probe shader timed on one host core on a small level beside it."""
    import time as _t
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import cubesynth
    import oracle_c
    out = {}
    n = 4096
    cs = C.c_void_p(stream.cuda_stream)
    gen = torch.Generator(device="cuda").manual_seed(11)
    nrm = torch.rand((n * n, 3), dtype=torch.float32, device="cuda", generator=gen)
    hkbcn.check(L.hkbcn_normalize_vectors_dev(0, nrm.data_ptr(), n * n, cs))
    packed = torch.empty(n * n * 2, dtype=torch.uint8, device="cuda")
    ms = _timed(torch, stream, flush, lambda: hkbcn.check(L.hkbcn_pack_normals_dev(0, 1, nrm.data_ptr(), packed.data_ptr(), n * n, cs)), 5, 3)
    gbs = n * n * 14 / (ms * 1e-3) / 1e9
    out["pack_normals_rg_bc5_4096"] = {"us": round(ms * 1e3, 1), "GB/s": round(gbs, 1), "frac": round(gbs / hbm, 4), "Gpix/s": round(n * n / ms / 1e6, 2)}
    ms = _timed(torch, stream, flush, lambda: hkbcn.check(L.hkbcn_normalize_vectors_dev(0, nrm.data_ptr(), n * n, cs)), 5, 3)
    gbs = n * n * 24 / (ms * 1e-3) / 1e9
    out["normalize_vectors_4096"] = {"us": round(ms * 1e3, 1), "GB/s": round(gbs, 1), "frac": round(gbs / hbm, 4)}
    mips = L.hkbcn_calc_num_mips(n, n)
    rough = torch.randint(0, 256, (L.hkbcn_chain_src_bytes(60, n, n, mips, 0),), dtype=torch.uint8, device="cuda", generator=gen)
    scratch = torch.empty(L.hkbcn_bake_scratch_bytes(n, n), dtype=torch.uint8, device="cuda")
    ms = _timed(torch, stream, flush, lambda: hkbcn.check(L.hkbcn_bake_roughness_dev(
        0, nrm.data_ptr(), rough.data_ptr(), scratch.data_ptr(), n, n, mips, 1, 2, 3, cs)), 3, 2)
    out["bake_roughness_vmf_4096_chain"] = {"ms": round(ms, 3), "Mpix/s": round(n * n / ms / 1e3, 1)}
    del nrm, packed, rough, scratch
    # environment probe: 6 x 512^2 source, the reference's maxLod 7 (128^2 .. 1^2) and 32^2 irradiance
    src, max_lod = 512, 7
    cube = torch.from_numpy(cubesynth.sky_cube(src, 1)).cuda()
    texels = L.hkbcn_envprobe_texels(max_lod)
    dst = torch.empty(texels * 4, dtype=torch.float32, device="cuda")
    ms = _timed(torch, stream, flush, lambda: hkbcn.check(L.hkbcn_envprobe_prefilter_dev(0, cube.data_ptr(), src, max_lod, dst.data_ptr(), cs)), 3, 2)
    out["envprobe_prefilter_512_lod7"] = {"ms": round(ms, 3), "Gsamples/s": round(texels * 1024 / ms / 1e6, 2), "texels": int(texels)}
    irr = torch.empty(6 * 32 * 32 * 4, dtype=torch.float32, device="cuda")
    ms_i = _timed(torch, stream, flush, lambda: hkbcn.check(L.hkbcn_irradiance_dev(0, cube.data_ptr(), src, 32, irr.data_ptr(), cs)), 3, 2)
    out["irradiance_512_to_32"] = {"ms": round(ms_i, 3), "Gsamples/s": round(6 * 32 * 32 * 252 * 63 / ms_i / 1e6, 2)}
    # the same shader arithmetic on one host core (oracle port), bounded: a 64^2 source and maxLod 3 (8^2 .. 1^2)
    small = cubesynth.sky_cube(64, 1)
    t0 = _t.perf_counter()
    oracle_c.envprobe_prefilter(small, 64, 3)
    dt = _t.perf_counter() - t0
    cpu_rate = 6 * (64 + 16 + 4 + 1) * 1024 / dt / 1e9
    out["envprobe_prefilter_cpu_port"] = {"Gsamples/s": round(cpu_rate, 5), "cores": 1, "sample": "64^2 source, maxLod 3 (510 texels x 1024 samples)",
                                          "kind": "port"}
    return out


def _profile_json():
    try:
        with open(os.path.join(ROOT, "profiles", "bc7_encode_ncu.json")) as f:
            return json.load(f)
    except (OSError, ValueError):
        return {}


def kernel_source_hash():
    """sha256 over the sources that make up bc7_encode_kernel: the committed ncu figures are valid for exactly this code"""
    import hashlib
    h = hashlib.sha256()
    for f in ("bc7_thread.cuh", "bc7_encode.cu", "hkbcn_common.cuh", "bc7_host_tables.h"):
        with open(os.path.join(ROOT, "hork-source_b200", "csrc", f), "rb") as fh:
            h.update(fh.read())
    return h.hexdigest()[:16]


def roofline_block(px, ms_step, clocks, hbm, which, hbm_achieved):
    """The governing roofline of bc7_encode_kernel is the SM issue rate, not HBM: achieved = warp instructions per second =
    (warp instructions per 4x4 block, counted by ncu on this very source, profiles/bc7_encode_ncu.json) x blocks / CUDA-event
    time of the step; peak = 148 SMs x 4 schedulers x the SM clock sampled during the timed region.  The instruction count is
    only valid for the source it was captured on: the profile carries a hash of the kernel sources, and a mismatch yields
    frac null + stale_profile instead of a silently stale fraction.  The HBM figures stay beside it."""
    prof = _profile_json()
    blocks = px // 16
    out = {"bound": "alu-issue", "kernel": "bc7_encode_kernel", "unit": "G warp-inst/s", "achieved": None, "peak": None, "frac": None,
           "hbm": {"achieved": hbm_achieved, "peak": hbm, "unit": "GB/s", "frac": hbm_achieved / hbm, "peak_source": which,
                   "algorithmic_bytes": 5 * px, "traffic": None},
           "hbm_frac": hbm_achieved / hbm, "traffic": None}
    stale = prof.get("source_hash") != kernel_source_hash()
    out["stale_profile"] = bool(stale)
    ipb = prof.get("warp_inst_per_block")
    if stale or not ipb or not clocks or not clocks.get("sm_mhz"):
        return out
    peak = 148 * 4 * clocks["sm_mhz"] * 1e6
    rate = ipb * blocks / (ms_step * 1e-3)
    traffic = prof["dram_bytes_per_px"] * px if prof.get("dram_bytes_per_px") else None
    out.update({"achieved": rate / 1e9, "peak": peak / 1e9, "frac": rate / peak, "warp_inst_per_block": ipb,
                "lanes_active": prof.get("lanes_active"), "pipe_alu_pct": prof.get("pipe_alu_pct"), "pipe_fmaheavy_pct": prof.get("pipe_fmaheavy_pct"),
                "issue_slots_busy_pct_ncu": prof.get("issue_slots_busy_pct"), "profile": prof.get("summary"), "traffic": traffic})
    out["hbm"]["traffic"] = traffic
    return out


def parity_block(levels, gpu_chain, ref_blocks):
    """in-run identity and PSNR: the GPU's blocks of the timed workload against the blocks the reference arm's encoder just
    produced for the same chain, both decoded with the REFERENCE decoder (bc7decomp) against the uncompressed source"""
    import refshim
    if ref_blocks is None:
        return None
    ref_chain = np.concatenate(ref_blocks)
    g = np.asarray(gpu_chain)
    same = (g.reshape(-1, 16) == ref_chain.reshape(-1, 16)).all(axis=1)

    def chain_psnr(chain):
        se, n, off = 0.0, 0, 0
        for l in levels:
            h, w, _ = l.shape
            nb = (w // 4) * (h // 4) * 16
            dec = refshim.decompress(BC7, chain[off:off + nb], w, h)
            d = dec.astype(np.int32) - l.ravel().astype(np.int32)
            se += float((d.astype(np.int64) ** 2).sum())
            n += d.size
            off += nb
        mse = se / n
        return 99.0 if mse == 0 else float(10 * np.log10(255.0 ** 2 / mse))
    pr = chain_psnr(ref_chain)
    pg = pr if same.all() else chain_psnr(g)
    return {"identical_blocks": int(same.sum()), "of": int(same.size), "psnr_gpu": round(pg, 4), "psnr_ref": round(pr, 4),
            "psnr": "RGBA, whole chain, decoded with the reference's bc7decomp"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--selftest-shard", action="store_true", help="CPU/gloo check of the N>1 plumbing (used by tests/)")
    args = ap.parse_args()
    if args.selftest_shard:
        run_selftest(args)
    elif args.impl == "reference":
        run_reference(args)
    else:
        run_gpu(args)


if __name__ == "__main__":
    main()
