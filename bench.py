#!/usr/bin/env python
"""bench.py -- the hot path's headline measurement.

Workload (BASELINE.json configs[4], the one the metric "at 1/2/4/8 B200" is quoted
on and the only one that shards): global paddle.sum of fp32 over axis 0, 8 GiB per
GPU ([524288, 4096] = 2^31 elements per rank), sharded by rows across N GPUs, local
reduce + NCCL allreduce of the [4096] partial (no collective at N = 1).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]
  torchrun ... bench.py --gpus N ...            (N > 1: one rank per GPU)

One JSON line on rank 0.  `value` = algorithmic GB/s of the whole job with inputs
resident in HBM; `e2e` = the same metric through the host-buffer entry point
(pinned host memory -> chunked H2D overlapped with the reduce -> result D2H);
`roofline` = the dominant kernel against the measured HBM peak; `cpu_baseline` =
the reference's own CPU arithmetic (oracle/_ref) on a bounded sample, 1 thread.
`--impl reference` times that CPU arithmetic on all host threads instead.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

ROWS, COLS = 524288, 4096                      # 8 GiB of fp32 per GPU
ALGO_BYTES = ROWS * COLS * 4 + COLS * 4        # 8,589,950,976 B per GPU per step
METRIC = "phi_global_sum_fp32_axis0_hbm_gbs"
UNIT = "GB/s"


def env_int(name, default):
    try:
        return int(os.environ.get(name, default))
    except ValueError:
        return default


class ClockSampler(threading.Thread):
    """SM clock + throttle reasons during the timed region, read in-process through NVML
    (the same counters `nvidia-smi --query-gpu=clocks.sm,clocks_event_reasons.*` prints;
    spawning nvidia-smi next to a 30 ms timed region perturbs it)."""

    def __init__(self, gpu_index, period=0.01):
        super().__init__(daemon=True)
        self.gpu = gpu_index
        self.period = period
        self.sm = []
        self.reasons = set()
        self.smax = None
        self.stop_flag = threading.Event()
        self.h = None
        try:
            import pynvml
            self.nv = pynvml
            pynvml.nvmlInit()
            # NVML indexes physical devices; honour CUDA_VISIBLE_DEVICES when it lists indices
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            idx = gpu_index
            if vis:
                try:
                    idx = int(vis.split(",")[gpu_index])
                except (ValueError, IndexError):
                    idx = gpu_index
            self.h = pynvml.nvmlDeviceGetHandleByIndex(idx)
            self.smax = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
        except Exception:
            self.h = None

    def run(self):
        if self.h is None:
            return
        nv = self.nv
        names = [("hw_slowdown", nv.nvmlClocksThrottleReasonHwSlowdown),
                 ("hw_thermal_slowdown", nv.nvmlClocksThrottleReasonHwThermalSlowdown),
                 ("sw_thermal_slowdown", nv.nvmlClocksThrottleReasonSwThermalSlowdown),
                 ("sw_power_cap", nv.nvmlClocksThrottleReasonSwPowerCap)]
        while not self.stop_flag.is_set():
            try:
                self.sm.append(float(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)))
                mask = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for n, bit in names:
                    if mask & bit:
                        self.reasons.add(n)
            except Exception:
                pass
            self.stop_flag.wait(self.period)

    def summary(self):
        return {"sm_mhz": statistics.median(self.sm) if self.sm else None, "sm_max_mhz": self.smax,
                "reasons": sorted(self.reasons), "samples": len(self.sm), "source": "nvml"}


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


def recorded_traffic():
    """dram bytes per launch from the committed ncu --set full capture, if one exists."""
    p = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(p):
        try:
            return json.load(open(p)).get("col_reduce_c5_dram_bytes_per_launch")
        except Exception:
            return None
    return None


# ------------------------------------------------------------------------------------ CPU arms
_CPU_BLOCK = None


def cpu_slab(sample_rows):
    """[sample_rows, COLS] fp32 uniform[0,1): a 4096-row random block tiled (cheap to make;
    the CPU arithmetic does not care that rows repeat)."""
    import numpy as np
    global _CPU_BLOCK
    if _CPU_BLOCK is None:
        _CPU_BLOCK = np.random.default_rng(123).random((4096, COLS), dtype=np.float32)
    x = np.empty((sample_rows, COLS), dtype=np.float32)
    for r0 in range(0, sample_rows, 4096):
        n = min(4096, sample_rows - r0)
        x[r0:r0 + n] = _CPU_BLOCK[:n]
    return x


def cpu_sum_axis0(threads, x):
    """The reference's CPU arithmetic for this workload on the slab x [rows, COLS].
    threads == 1: phi::funcs::SumFunctor on one Eigen::DefaultDevice call, exactly what
    phi::SumKernel<float, CPUContext> runs.  threads > 1: row blocks on separate host
    threads (each the same call), partials folded by one more call.  Returns (seconds, kind)."""
    import ctypes as C
    import numpy as np
    from oracle import pyoracle as O
    kind = "reference" if O.have_ref() else "port"
    sample_rows = x.shape[0]
    out = np.empty(COLS, dtype=np.float32)

    def one(xb, ob):
        if kind == "reference":
            dims = (C.c_int64 * 9)(*xb.shape)
            ax = (C.c_int64 * 9)(0)
            rc = O.ref().phi_ref_sum_float(xb.ctypes.data_as(C.c_void_p), 2, dims, 1, ax, 0,
                                           ob.ctypes.data_as(C.c_void_p))
            assert rc == 0
        else:
            ob[...] = O.reduce("sum", xb, [0])

    t0 = time.perf_counter()
    if threads <= 1:
        one(x, out)
    else:
        parts = np.empty((threads, COLS), dtype=np.float32)
        bounds = [sample_rows * t // threads for t in range(threads + 1)]
        ths = [threading.Thread(target=one, args=(x[bounds[t]:bounds[t + 1]], parts[t]))
               for t in range(threads)]
        for t in ths:
            t.start()
        for t in ths:
            t.join()
        one(parts, out)
    dt = time.perf_counter() - t0
    chk = x[:, :64].astype(np.float64).sum(0)
    np.testing.assert_allclose(out[:64], chk, rtol=1e-4)
    return dt, kind


def run_reference_arm(args):
    rank = env_int("RANK", 0)
    if rank != 0:
        return  # rank 0 alone runs the CPU arm
    threads = os.cpu_count() or 1
    # size the slab from a probe that is already far larger than the CPU caches, so that the
    # whole --steps/--warmup run stays within a few minutes
    probe = cpu_slab(32768)                                   # 512 MiB
    t_probe, kind = cpu_sum_axis0(threads, probe)
    per_row = max(t_probe / probe.shape[0], 1e-9)
    budget_s = min(150.0 / max(args.steps + args.warmup, 1), 6.0)
    sample_rows = int(min(ROWS // 4, max(32768, budget_s / per_row)))
    sample_rows -= sample_rows % threads
    x = probe if sample_rows == probe.shape[0] else cpu_slab(sample_rows)
    for _ in range(args.warmup):
        cpu_sum_axis0(threads, x)
    times = [cpu_sum_axis0(threads, x)[0] for _ in range(args.steps)]
    t = sum(times) / len(times)
    sample_bytes = sample_rows * COLS * 4 + COLS * 4
    gbs = sample_bytes / t / 1e9
    line = {
        "impl": "reference", "metric": METRIC, "value": gbs, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": t * 1e3 * (ALGO_BYTES / sample_bytes),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic",
        "config": {"workload": "C5 global paddle.sum fp32 axis=0, [524288,4096] per GPU (8 GiB)",
                   "note": "reference CPU arithmetic timed on a bounded row slab; ms_per_step is the "
                           "slab time scaled to one full 8 GiB shard"},
        "cpu_baseline": {"value": gbs, "unit": UNIT, "cores": threads, "kind": kind,
                         "sample": f"[{sample_rows},{COLS}] fp32 slab ({sample_bytes / 2**20:.0f} MiB) per step, "
                                   f"row blocks on {threads} host threads"},
        "e2e": {"value": gbs, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------ GPU arm
def fill_shard(torch, x, rank):
    """uniform [0,1) keyed by the global row index (counter-based: same on every run)."""
    g = torch.Generator(device=x.device)
    chunk = 32768
    for r0 in range(0, x.shape[0], chunk):
        g.manual_seed(123 + rank * 100003 + r0)
        x[r0:r0 + chunk].uniform_(0.0, 1.0, generator=g)


def time_steps(torch, fn, steps, warmup, dist=None):
    for _ in range(warmup):
        fn()
    if dist is not None:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    if dist is not None:
        dist.barrier()
    return e0.elapsed_time(e1) / 1e3  # seconds for `steps` steps


def other_configs(torch, pb):
    """Kernel-only GB/s of C1-C4 (parity cases, reported for context; inputs > L2).
    Descriptors and outputs are prepared once and the C ABI is called back to back, so the
    figure is the kernel's, not the Python wrapper's."""
    import ctypes as C
    from paddle2_b200 import _lib as L
    lib = L.lib()
    res = {}

    def t(fn, nbytes, iters=20, warm=5):
        """`iters` back-to-back launches replayed from a CUDA graph: at 40-100 us per kernel the
        Python/ctypes call (~30 us) would otherwise be what is measured."""
        for _ in range(warm):
            fn()
        torch.cuda.synchronize()
        try:
            side = torch.cuda.Stream()
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.stream(side):
                nonlocal_cur.value = side.cuda_stream
                with torch.cuda.graph(graph, stream=side):
                    for _ in range(iters):
                        fn()
            nonlocal_cur.value = torch.cuda.current_stream().cuda_stream
            graph.replay()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            graph.replay()
            e1.record()
            torch.cuda.synchronize()
            s = e0.elapsed_time(e1) / 1e3
        except Exception:
            nonlocal_cur.value = torch.cuda.current_stream().cuda_stream
            torch.cuda.synchronize()
            s = time_steps(torch, fn, iters, 0)
        return round(nbytes / (s / iters) / 1e9, 1)

    nonlocal_cur = C.c_void_p(torch.cuda.current_stream().cuda_stream)

    def stream():
        return nonlocal_cur

    x = torch.rand(32, 1024, 4096, device="cuda")
    out = torch.empty(32, 1024, device="cuda")
    d = L.ReduceDesc()
    d.op, d.in_dtype, d.out_dtype, d.rank = L.SUM, L.FLOAT32, L.FLOAT32, 3
    d.dims[0], d.dims[1], d.dims[2] = 32, 1024, 4096
    d.naxes = 1
    d.axes[0] = 2
    d.x, d.out = x.data_ptr(), out.data_ptr()
    res["C1_sum_fp32_32x1024x4096_lastaxis"] = t(lambda: L.check(lib.b200phi_reduce(C.byref(d), stream())), 537001984)
    assert torch.allclose(out, x.sum(-1), rtol=1e-4)
    del x, out

    x = torch.rand(256, 64, 112, 112, device="cuda") * 2 - 1
    y = torch.rand(64, 1, 1, device="cuda") * 2 - 1
    out = torch.empty_like(x)
    e = L.ElementwiseDesc()
    e.binary, e.act, e.compound, e.dtype, e.arity, e.axis = L.ADD, L.ACT_RELU, 0, L.FLOAT32, 2, -1
    for i, tt in enumerate((x, y)):
        e.in_[i].rank = tt.dim()
        for j, sz in enumerate(tt.shape):
            e.in_[i].dims[j] = sz
        e.in_[i].data = tt.data_ptr()
    e.out_rank = 4
    for j, sz in enumerate(x.shape):
        e.out_dims[j] = sz
    e.out = out.data_ptr()
    res["C2_add_relu_fp32_256x64x112x112"] = t(lambda: L.check(lib.b200phi_elementwise(C.byref(e), stream())), 1644167424)
    assert torch.equal(out[3], torch.relu(x[3] + y))
    del x, y, out

    def arg_desc(xt, idx, val):
        a = L.ArgReduceDesc()
        a.op, a.in_dtype, a.index_dtype, a.rank = L.ARGMAX, pb.ops._phi_dtype(xt.dtype), L.INT64, 2
        a.dims[0], a.dims[1] = xt.shape
        a.axis = 1
        a.x, a.out, a.max_out = xt.data_ptr(), idx.data_ptr(), val.data_ptr()
        return a

    x = torch.randint(-(2 ** 40), 2 ** 40, (8192, 32000), device="cuda", dtype=torch.int64)
    idx = torch.empty(8192, device="cuda", dtype=torch.int64)
    val = torch.empty(8192, device="cuda", dtype=torch.int64)
    a = arg_desc(x, idx, val)
    res["C3a_argmax+max_int64_8192x32000"] = t(lambda: L.check(lib.b200phi_argreduce(C.byref(a), stream())), 2097283072)
    assert torch.equal(val, x.max(1).values)
    del x
    x = torch.randn(8192, 32000, device="cuda").bfloat16()
    val = torch.empty(8192, device="cuda", dtype=torch.bfloat16)
    a = arg_desc(x, idx, val)
    res["C3b_argmax+max_bf16_8192x32000"] = t(lambda: L.check(lib.b200phi_argreduce(C.byref(a), stream())), 524369920)
    assert torch.equal(val, x.max(1).values)
    del x

    x = torch.randn(16384, 8192, device="cuda").bfloat16()
    m = torch.empty(16384, device="cuda", dtype=torch.bfloat16)
    v = torch.empty(16384, device="cuda", dtype=torch.bfloat16)
    r = L.RowStatsDesc()
    r.in_dtype, r.out_dtype, r.rows, r.cols, r.unbiased = L.BFLOAT16, L.BFLOAT16, 16384, 8192, 1
    r.x, r.mean, r.var = x.data_ptr(), m.data_ptr(), v.data_ptr()
    res["C4_row_mean_var_bf16_16384x8192"] = t(lambda: L.check(lib.b200phi_rowstats(C.byref(r), stream())), 268500992)
    assert torch.allclose(v.float(), x.float().var(1), rtol=2e-2)
    del x
    torch.cuda.empty_cache()

    # SURVEY 8(f) rows 1-2 (context only): row logsumexp, max_grad, a compare
    x = torch.randn(8192, 32000, device="cuda").bfloat16()
    o = torch.empty(8192, device="cuda", dtype=torch.bfloat16)
    ld = L.LogsumexpDesc()
    ld.dtype, ld.rows, ld.cols, ld.x, ld.out = L.BFLOAT16, 8192, 32000, x.data_ptr(), o.data_ptr()
    res["F_logsumexp_bf16_8192x32000"] = t(lambda: L.check(lib.b200phi_logsumexp(C.byref(ld), stream())), 524304384)
    assert torch.allclose(o.float(), torch.logsumexp(x.float(), 1), rtol=1e-2)

    def map_desc(op, ins, outt):
        md = L.MapDesc()
        md.op, md.in_dtype, md.out_dtype, md.arity, md.axis = L.MAP_OPS[op], pb.ops._phi_dtype(ins[0].dtype), 0, len(ins), -1
        for i, tt in enumerate(ins):
            md.in_[i].rank = tt.dim()
            for j, sz in enumerate(tt.shape):
                md.in_[i].dims[j] = sz
            md.in_[i].data = tt.data_ptr()
        md.out_rank = outt.dim()
        for j, sz in enumerate(outt.shape):
            md.out_dims[j] = sz
        md.out = outt.data_ptr()
        return md

    mx = x.max(1, keepdim=True).values
    g = torch.randn(8192, 1, device="cuda").bfloat16()
    dx = torch.empty_like(x)
    md = map_desc("equal_mul", [x, mx, g], dx)
    res["F_max_grad_bf16_8192x32000"] = t(lambda: L.check(lib.b200phi_map(C.byref(md), stream())), 1048576000)
    assert torch.equal(dx, (x == mx).to(x.dtype) * g)
    del x, dx
    torch.cuda.empty_cache()
    x = torch.rand(256, 64, 112, 112, device="cuda")
    y = torch.rand(64, 1, 1, device="cuda")
    b = torch.empty(x.shape, device="cuda", dtype=torch.bool)
    md2 = map_desc("less_than", [x, y], b)
    res["F_less_than_fp32_256x64x112x112_vs_64x1x1"] = t(lambda: L.check(lib.b200phi_map(C.byref(md2), stream())), 1027604480)
    assert torch.equal(b, x < y)
    del x, y, b
    torch.cuda.empty_cache()
    return res


def run_gpu_arm(args):
    import torch
    import torch.distributed as dist
    import paddle2_b200 as pb
    from paddle2_b200 import _lib as L
    from paddle2_b200.dist import ShardedReducer
    from paddle2_b200.host import HostAxis0Reducer

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: there is no CPU fallback for the product path")
    L.lib()  # fail loudly if the extension is missing
    world = env_int("WORLD_SIZE", 1)
    rank = env_int("RANK", 0)
    local_rank = env_int("LOCAL_RANK", 0)
    if world != args.gpus and world > 1:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}")
    torch.cuda.set_device(local_rank)
    multi = world > 1
    if multi:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    reducer = ShardedReducer() if multi else None

    x = torch.empty(ROWS, COLS, device="cuda", dtype=torch.float32)
    fill_shard(torch, x, rank)
    out = torch.empty(COLS, device="cuda", dtype=torch.float32)

    # pre-size the scratch once (dev_ctx.Alloc would hand the same block back every call)
    import ctypes as C
    d = L.ReduceDesc()
    d.op, d.in_dtype, d.out_dtype, d.rank = L.SUM, L.FLOAT32, L.FLOAT32, 2
    d.dims[0], d.dims[1] = ROWS, COLS
    d.naxes = 1
    d.axes[0] = 0
    d.x, d.out = x.data_ptr(), out.data_ptr()
    lib = L.lib()
    ws_bytes = lib.b200phi_reduce_workspace(C.byref(d))
    ws = torch.empty(max(ws_bytes, 256), dtype=torch.uint8, device="cuda")
    d.workspace, d.workspace_bytes = ws.data_ptr(), ws.numel()
    comm = reducer.comm if reducer is not None else None

    # N > 1: ncclAllReduce on the reduce's stream by default.  B200PHI_COLLECTIVE=p2p selects the
    # fused kernel (local reduce + exchange over NVLink peer memory in one launch); measured at
    # N=2 it is 1.212 ms/step against 1.204 for the NCCL route (profiles/r01_multigpu.md), so
    # the faster one is the default
    p2p = None
    collective = "none (single GPU)"
    if multi:
        collective = "ncclAllReduce after the local reduce (b200phi_sharded_reduce)"
        if os.environ.get("B200PHI_COLLECTIVE", "nccl") == "p2p":
            try:
                from paddle2_b200.dist import P2PReducer
                p2p = P2PReducer(max_out_numel=COLS, max_tiles=64)
                collective = "fused in the reduce kernel over NVLink peer memory (b200phi_sharded_reduce_p2p)"
            except Exception as exc:  # e.g. CUDA IPC not permitted in this container
                p2p = None
                collective += f" [p2p unavailable: {str(exc)[:80]}]"
        ok = torch.tensor([1 if p2p is not None else 0], device="cuda")
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)
        if int(ok.item()) == 0:
            p2p = None
    state = {"epoch": 0}

    def step():
        stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
        if p2p is not None:
            p2p.epoch += 1
            L.check(lib.b200phi_sharded_reduce_p2p(C.byref(d), 0, C.byref(p2p.g), p2p.epoch, stream))
        else:
            L.check(lib.b200phi_sharded_reduce(C.byref(d), 0, comm, stream))

    sampler = ClockSampler(local_rank) if rank == 0 else None
    if sampler:
        sampler.start()
    secs = time_steps(torch, step, args.steps, args.warmup, dist if multi else None)
    if sampler:
        sampler.stop_flag.set()
        sampler.join(timeout=3)
    t = torch.tensor([secs], device="cuda", dtype=torch.float64)
    if multi:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    secs = float(t.item())
    ms_per_step = secs / args.steps * 1e3
    value = ALGO_BYTES * world / (secs / args.steps) / 1e9

    if p2p is not None:  # the fused exchange must agree with the NCCL route on the same shards
        ref = reducer.sum(x)
        torch.cuda.synchronize()
        assert torch.allclose(out, ref, rtol=1e-6), "fused p2p result differs from the NCCL route"
    # correctness of what was timed: column sums against an fp64 torch reduction of a slab + full
    chk = x[:8192].double().sum(0)
    got_slab = pb.sum(x[:8192], axis=0).double()
    assert torch.allclose(got_slab, chk, rtol=1e-5), "bench result check failed"

    # ---- end to end through the host-buffer entry point ------------------------------------
    e2e = None
    try:
        import psutil
        avail = psutil.virtual_memory().available
    except Exception:
        avail = 0
    shard_bytes = ROWS * COLS * 4
    host_rows = ROWS if (not multi and avail > 4 * shard_bytes) else 65536   # else: 1 GiB staging slab
    xh = torch.empty(host_rows, COLS, dtype=torch.float32, pin_memory=True)
    xh.copy_(x[:host_rows])
    hred = HostAxis0Reducer(COLS, torch.float32, "cuda", chunk_rows=16384)
    out_h = torch.empty(COLS, dtype=torch.float32, pin_memory=True)
    reps = ROWS // host_rows          # a smaller staging slab is streamed `reps` times per step
    h2d = d2h = 0

    def e2e_step():
        nonlocal h2d, d2h
        h2d = d2h = 0
        for _ in range(reps):
            hred(xh, out_h)
            h2d += hred.h2d_bytes
            d2h += hred.d2h_bytes
        if multi:   # the [C] partial is combined exactly as in the resident path
            out.copy_(out_h, non_blocking=True)
            reducer.reduce(L.SUM, out.view(1, COLS), out=out)
            out_h.copy_(out, non_blocking=True)
            torch.cuda.current_stream().synchronize()

    e2e_steps = min(args.steps, 5)
    e2e_step()
    if multi:
        dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        e2e_step()
    torch.cuda.synchronize()
    te = (time.perf_counter() - t0) / e2e_steps
    tt = torch.tensor([te], device="cuda", dtype=torch.float64)
    if multi:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    te = float(tt.item())
    if not multi:
        assert torch.allclose(out_h.double(), x[:host_rows].double().sum(0).cpu(), rtol=1e-4), \
            "e2e result check failed"
    e2e = {"value": ALGO_BYTES * world / te / 1e9, "unit": UNIT, "h2d_bytes_per_step": int(h2d),
           "d2h_bytes_per_step": int(d2h), "steps": e2e_steps,
           "host_buffer": f"[{host_rows},{COLS}] pinned fp32, streamed {reps}x per step in 16384-row chunks"}
    del xh

    if p2p is not None:
        p2p.close()
    if rank != 0:
        if multi:
            dist.barrier()
            dist.destroy_process_group()
        return

    peak, peak_src = measured_peak()
    kernel_gbs = ALGO_BYTES / (secs / args.steps) / 1e9 if not multi else None
    roofline = {"bound": "hbm", "kernel": "col_reduce_kernel<float,float,SumOp,4,8>",
                "achieved": kernel_gbs if kernel_gbs is not None else value / world,
                "peak": peak, "unit": UNIT, "peak_source": peak_src,
                "frac": (kernel_gbs if kernel_gbs is not None else value / world) / peak,
                "frac_of_nominal_8000": (kernel_gbs if kernel_gbs is not None else value / world) / 8000.0,
                "traffic": recorded_traffic(),
                "algorithmic_bytes_per_launch": ALGO_BYTES,
                "note": "CUDA events over the timed region on the launching stream; at N>1 the step also "
                        "holds the 16 KiB allreduce"}
    cpu_baseline = None
    others = None
    if not multi:
        t_cpu, kind = cpu_sum_axis0(1, cpu_slab(16384))
        per_row = t_cpu / 16384
        rows_s = int(min(ROWS // 8, max(16384, 12.0 / per_row)))
        t_cpu, kind = cpu_sum_axis0(1, cpu_slab(rows_s))
        sb = rows_s * COLS * 4 + COLS * 4
        cpu_baseline = {"value": sb / t_cpu / 1e9, "unit": UNIT, "cores": 1, "kind": kind,
                        "sample": f"[{rows_s},{COLS}] fp32 slab of the same shard ({sb / 2**20:.0f} MiB), "
                                  f"one phi::funcs::SumFunctor call on Eigen::DefaultDevice, {t_cpu:.2f} s"}
        if not args.no_others:
            del x
            torch.cuda.empty_cache()
            others = other_configs(torch, pb)
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "C5 global paddle.sum fp32 axis=0, [524288,4096] per GPU (8 GiB), "
                               "row-sharded, local reduce + ncclAllReduce of the [4096] partial",
                   "l2": "inputs (8 GiB per GPU) are far larger than the 126 MB L2; no flush needed",
                   "parallelism": f"rows sharded over {world} GPU(s), one process per GPU",
                   "collective": collective},
        "clocks": sampler.summary() if sampler else None,
        "e2e": e2e, "gpu_launches": args.steps,
        "roofline": roofline, "cpu_baseline": cpu_baseline,
    }
    if others:
        line["other_configs_gbs"] = others
    print(json.dumps(line), flush=True)
    if multi:
        dist.barrier()
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-others", action="store_true", help="skip the C1-C4 context numbers")
    args = ap.parse_args()
    if args.warmup < 3:
        args.warmup = 3
    if args.impl == "reference":
        run_reference_arm(args)
    else:
        run_gpu_arm(args)


if __name__ == "__main__":
    main()
