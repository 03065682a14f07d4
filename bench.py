#!/usr/bin/env python
"""bench.py -- the hot path's headline measurement.

Workload (BASELINE.json configs[4], the one the metric "at 1/2/4/8 B200" is quoted
on and the only one that shards): global paddle.sum of fp32 over axis 0, 8 GiB per
GPU ([524288, 4096] = 2^31 elements per rank), sharded by rows across N GPUs, local
reduce + NCCL allreduce of the [4096] partial (no collective at N = 1).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]
  torchrun ... bench.py --gpus N ...            (N > 1: one rank per GPU)

Inputs are the closed-form generator x[r, c] = ((131 r + 7 c) mod 1024) / 1024 keyed by the
GLOBAL row index (uniform over multiples of 1/1024 in [0, 1); every 1024 consecutive rows sum to
511.5 per column), so the all-reduced result is known exactly on every rank without a host copy
and is verified after the timed region (`result_check`).

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
DEFAULT_COLLECTIVE = "p2p"
WORKLOAD = "C5 global paddle.sum fp32 axis=0, [524288,4096] per GPU (8 GiB)"   # both arms


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


def recorded_traffic(key="col_reduce_c5_dram_bytes_per_launch"):
    """dram bytes per launch from the committed ncu --set full capture (profiles/traffic.json):
    a recorded constant of that capture, not something measured in this run."""
    p = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(p):
        try:
            return json.load(open(p)).get(key)
        except Exception:
            return None
    return None


# ------------------------------------------------------------------------------------ CPU arms
_CPU_BLOCK = None


def closed_form_block_np():
    """rows 0..1023 of the closed-form tensor (it repeats with period 1024 rows)."""
    import numpy as np
    global _CPU_BLOCK
    if _CPU_BLOCK is None:
        r = np.arange(1024, dtype=np.int64)[:, None] * 131
        c = np.arange(COLS, dtype=np.int64)[None, :] * 7
        _CPU_BLOCK = (((r + c) % 1024).astype(np.float32) / np.float32(1024.0))
    return _CPU_BLOCK


def cpu_slab(sample_rows):
    """rows [0, sample_rows) of rank 0's shard: the same values the GPU arm reduces."""
    import numpy as np
    blk = closed_form_block_np()
    x = np.empty((sample_rows, COLS), dtype=np.float32)
    for r0 in range(0, sample_rows, 1024):
        n = min(1024, sample_rows - r0)
        x[r0:r0 + n] = blk[:n]
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
    # the reference adds a column's rows one after another in fp32: its own error grows with the
    # row count (5e-4 at 65536 rows), hence the width of this sanity bound
    chk = x[:, :64].astype(np.float64).sum(0)
    np.testing.assert_allclose(out[:64], chk, rtol=5e-3)
    return dt, kind


def run_reference_arm(args):
    rank = env_int("RANK", 0)
    if rank != 0:
        return  # rank 0 alone runs the CPU arm
    threads = os.cpu_count() or 1
    # One step = the whole [524288, 4096] shard when the host can do it within the run budget
    # (about 2.5 s on 16 threads); otherwise a row slab sized from a probe, and it says so.
    probe = cpu_slab(32768)                                   # 512 MiB
    t_probe, kind = cpu_sum_axis0(threads, probe)
    per_row = max(t_probe / probe.shape[0], 1e-9)
    budget_s = 240.0 / max(args.steps + args.warmup, 1)
    try:
        import psutil
        ram_rows = int(psutil.virtual_memory().available * 0.6) // (COLS * 4)
    except Exception:
        ram_rows = ROWS // 4
    sample_rows = int(min(ROWS, ram_rows, max(32768, budget_s / per_row)))
    if sample_rows < ROWS:
        sample_rows -= sample_rows % (1024 * threads) if sample_rows >= 1024 * threads else 0
    full = sample_rows == ROWS
    x = probe if sample_rows == probe.shape[0] else cpu_slab(sample_rows)
    for _ in range(args.warmup):
        cpu_sum_axis0(threads, x)
    times = [cpu_sum_axis0(threads, x)[0] for _ in range(args.steps)]
    t = sum(times) / len(times)
    sample_bytes = sample_rows * COLS * 4 + COLS * 4
    gbs = sample_bytes / t / 1e9
    sample = (f"the full [{ROWS},{COLS}] fp32 shard per step" if full else
              f"[{sample_rows},{COLS}] fp32 slab ({sample_bytes / 2**20:.0f} MiB) of the shard per step; "
              f"ms_per_step is the slab time scaled to the full shard")
    line = {
        "impl": "reference", "metric": METRIC, "value": gbs, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": t * 1e3 * (ALGO_BYTES / sample_bytes),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic",
        "config": {"workload": WORKLOAD,
                   "note": "the reference's CPU arithmetic (phi::funcs::SumFunctor on Eigen::DefaultDevice), "
                           "row blocks on all host threads; " + sample},
        "cpu_baseline": {"value": gbs, "unit": UNIT, "cores": threads, "kind": kind,
                         "sample": sample + f", row blocks on {threads} host threads"},
        "e2e": {"value": gbs, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------ GPU arm
def fill_shard(torch, x, rank):
    """x[r, c] = ((131 (rank*ROWS + r) + 7 c) mod 1024) / 1024 -- keyed by the global row."""
    cols = x.shape[1]
    c = torch.arange(cols, device=x.device, dtype=torch.int64)[None, :] * 7
    chunk = 16384
    row0 = rank * x.shape[0]
    for r0 in range(0, x.shape[0], chunk):
        r = torch.arange(row0 + r0, row0 + min(r0 + chunk, x.shape[0]), device=x.device,
                         dtype=torch.int64)[:, None] * 131
        x[r0:r0 + r.shape[0]] = ((r + c) % 1024).float() / 1024.0


def expected_colsum(total_rows):
    """Exact column sum of the closed-form tensor over `total_rows` rows (a multiple of 1024)."""
    assert total_rows % 1024 == 0
    return (total_rows // 1024) * 511.5


def bind_to_gpu_numa(torch, local_rank):
    """Best effort: run this process on the CPUs of the GPU's NUMA node, so that the pinned host
    buffer allocated next is first-touched (and pinned) in node-local memory."""
    try:
        bus = torch.cuda.get_device_properties(local_rank).pci_bus_id
        dom = torch.cuda.get_device_properties(local_rank).pci_domain_id
        dev = torch.cuda.get_device_properties(local_rank).pci_device_id
        path = f"/sys/bus/pci/devices/{dom:04x}:{bus:02x}:{dev:02x}.0/numa_node"
        node = int(open(path).read().strip())
        if node < 0:
            return None
        cpus = set()
        for part in open(f"/sys/devices/system/node/node{node}/cpulist").read().strip().split(","):
            lo, _, hi = part.partition("-")
            cpus.update(range(int(lo), int(hi or lo) + 1))
        cpus &= os.sched_getaffinity(0)
        if cpus:
            os.sched_setaffinity(0, cpus)
            return node
    except Exception:
        return None
    return None


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


ROT = 3   # input buffers rotated per config, so a replay never finds its input in the 126 MB L2


def cpu_baselines_c1_c4():
    """The reference's CPU arithmetic (oracle/_ref: its own functors + Eigen, 1 thread -- phi CPU
    kernels run on Eigen::DefaultDevice) for C1-C4, each on a stated slab of the same shapes.
    bf16 inputs are widened to fp32 first: phi registers no bf16 CPU kernel for max / argmax /
    mean (cpu/reduce_max_kernel.cc:38-39, cpu/arg_min_max_kernel.cc:200-224)."""
    import numpy as np
    from oracle import pyoracle as O
    impl = "ref" if O.have_ref() else "port"
    kind = "reference" if impl == "ref" else "port"
    rng = np.random.default_rng(123)
    res = {}

    def rec(name, nbytes, sample, fn):
        t0 = time.perf_counter()
        fn()
        dt = time.perf_counter() - t0
        res[name] = {"value": round(nbytes / dt / 1e9, 3), "unit": UNIT, "cores": 1, "kind": kind,
                     "sample": f"{sample}, {dt:.2f} s"}

    x = rng.random((32, 1024, 4096), dtype=np.float32)
    rec("C1_sum_fp32_32x1024x4096_lastaxis", x.nbytes + 32 * 1024 * 4, "the full [32,1024,4096] fp32 tensor",
        lambda: O.reduce("sum", x, [-1], impl=impl))
    x = rng.random((32, 64, 112, 112), dtype=np.float32) * 2 - 1
    y = rng.random((64, 1, 1), dtype=np.float32) * 2 - 1
    rec("C2_add_relu_fp32_256x64x112x112", 2 * x.nbytes + 256, "[32,64,112,112] slab (1/8 of the batch)",
        lambda: O.elementwise("add", [x, y], act="relu", impl=impl))
    x = rng.integers(-(2 ** 40), 2 ** 40, (2048, 32000), dtype=np.int64)

    def c3(xx):
        O.argreduce("argmax", xx, -1, impl=impl)
        O.reduce("max", xx, [-1], impl=impl)

    rec("C3a_argmax+max_int64_8192x32000", 2 * x.nbytes + 2048 * 16, "[2048,32000] int64 slab, argmax then max "
        "(two passes: the CPU has no fused form)", lambda: c3(x))
    xf = rng.standard_normal((2048, 32000)).astype(np.float32)
    rec("C3b_argmax+max_bf16_8192x32000", 2 * xf.nbytes + 2048 * 12, "[2048,32000] slab widened to fp32 "
        "(4 B/element read on the CPU), argmax then max", lambda: c3(xf))
    xf = (rng.standard_normal((4096, 8192)) * 2 ** 0.5 + 0.5).astype(np.float32)

    def c4():
        O.reduce("mean", xf, [-1], impl=impl)
        O.var_composite(xf, -1, impl=impl)

    rec("C4_row_mean_var_bf16_16384x8192", xf.nbytes // 2 + 4096 * 4, "[4096,8192] slab widened to fp32; rate "
        "counted in the bf16 tensor's algorithmic bytes; mean, then paddle.var's composite "
        "(mean, subtract, multiply, sum, divide: stat.py:184-197)", c4)
    return res


def other_configs(torch, pb):
    """Kernel-only GB/s of C1-C4 (parity cases, reported for context).  Descriptors and outputs
    are prepared once and the C ABI is called back to back, so the figure is the kernel's, not the
    Python wrapper's; ROT input buffers rotate so that no launch can find its input in L2."""
    import ctypes as C
    from paddle2_b200 import _lib as L
    lib = L.lib()
    res = {}

    def t(fns, nbytes, iters=21, warm=6):
        """`iters` back-to-back launches (round-robin over fns) replayed from a CUDA graph: at
        40-100 us per kernel the Python/ctypes call (~30 us) would otherwise be what is measured."""
        for i in range(warm):
            fns[i % len(fns)]()
        torch.cuda.synchronize()
        try:
            side = torch.cuda.Stream()
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.stream(side):
                nonlocal_cur.value = side.cuda_stream
                with torch.cuda.graph(graph, stream=side):
                    for i in range(iters):
                        fns[i % len(fns)]()
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
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for i in range(iters):
                fns[i % len(fns)]()
            e1.record()
            torch.cuda.synchronize()
            s = e0.elapsed_time(e1) / 1e3
        return round(nbytes / (s / iters) / 1e9, 1)

    nonlocal_cur = C.c_void_p(torch.cuda.current_stream().cuda_stream)

    def stream():
        return nonlocal_cur

    def reduce_call(d):
        return lambda: L.check(lib.b200phi_reduce(C.byref(d), stream()))

    xs = [torch.rand(32, 1024, 4096, device="cuda") for _ in range(ROT)]
    out = torch.empty(32, 1024, device="cuda")
    ds = []
    for x in xs:
        d = L.ReduceDesc()
        d.op, d.in_dtype, d.out_dtype, d.rank = L.SUM, L.FLOAT32, L.FLOAT32, 3
        d.dims[0], d.dims[1], d.dims[2] = 32, 1024, 4096
        d.naxes = 1
        d.axes[0] = 2
        d.x, d.out = x.data_ptr(), out.data_ptr()
        ds.append(d)
    res["C1_sum_fp32_32x1024x4096_lastaxis"] = t([reduce_call(d) for d in ds], 537001984)
    assert torch.allclose(out, xs[(21 - 1) % ROT].sum(-1), rtol=1e-4)
    del xs, out

    xs = [torch.rand(256, 64, 112, 112, device="cuda") * 2 - 1 for _ in range(2)]
    y = torch.rand(64, 1, 1, device="cuda") * 2 - 1
    outs = [torch.empty_like(xs[0]) for _ in range(2)]
    es = []
    for x, out in zip(xs, outs):
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
        es.append(e)
    res["C2_add_relu_fp32_256x64x112x112"] = t(
        [(lambda e=e: L.check(lib.b200phi_elementwise(C.byref(e), stream()))) for e in es], 1644167424)
    assert torch.equal(outs[0][3], torch.relu(xs[0][3] + y))
    del xs, y, outs

    def arg_desc(xt, idx, val):
        a = L.ArgReduceDesc()
        a.op, a.in_dtype, a.index_dtype, a.rank = L.ARGMAX, pb.ops._phi_dtype(xt.dtype), L.INT64, 2
        a.dims[0], a.dims[1] = xt.shape
        a.axis = 1
        a.x, a.out, a.max_out = xt.data_ptr(), idx.data_ptr(), val.data_ptr()
        return a

    def arg_call(a):
        return lambda: L.check(lib.b200phi_argreduce(C.byref(a), stream()))

    xs = [torch.randint(-(2 ** 40), 2 ** 40, (8192, 32000), device="cuda", dtype=torch.int64) for _ in range(2)]
    idx = torch.empty(8192, device="cuda", dtype=torch.int64)
    val = torch.empty(8192, device="cuda", dtype=torch.int64)
    descs = [arg_desc(x, idx, val) for x in xs]
    res["C3a_argmax+max_int64_8192x32000"] = t([arg_call(a) for a in descs], 2097283072, iters=20)
    assert torch.equal(val, xs[1].max(1).values)
    del xs
    xs = [torch.randn(8192, 32000, device="cuda").bfloat16() for _ in range(ROT)]
    val = torch.empty(8192, device="cuda", dtype=torch.bfloat16)
    descs = [arg_desc(x, idx, val) for x in xs]
    res["C3b_argmax+max_bf16_8192x32000"] = t([arg_call(a) for a in descs], 524369920)
    assert torch.equal(val, xs[(21 - 1) % ROT].max(1).values)
    del xs

    xs = [torch.randn(16384, 8192, device="cuda").bfloat16() for _ in range(ROT)]
    m = torch.empty(16384, device="cuda", dtype=torch.bfloat16)
    v = torch.empty(16384, device="cuda", dtype=torch.bfloat16)
    rs = []
    for x in xs:
        r = L.RowStatsDesc()
        r.in_dtype, r.out_dtype, r.rows, r.cols, r.unbiased = L.BFLOAT16, L.BFLOAT16, 16384, 8192, 1
        r.x, r.mean, r.var = x.data_ptr(), m.data_ptr(), v.data_ptr()
        rs.append(r)
    res["C4_row_mean_var_bf16_16384x8192"] = t(
        [(lambda r=r: L.check(lib.b200phi_rowstats(C.byref(r), stream()))) for r in rs], 268500992)
    assert torch.allclose(v.float(), xs[(21 - 1) % ROT].float().var(1), rtol=2e-2)
    del xs
    torch.cuda.empty_cache()

    # SURVEY 8(f) rows 1-2 (context only): row logsumexp, max_grad, a compare
    xs = [torch.randn(8192, 32000, device="cuda").bfloat16() for _ in range(ROT)]
    o = torch.empty(8192, device="cuda", dtype=torch.bfloat16)
    lds = []
    for x in xs:
        ld = L.LogsumexpDesc()
        ld.dtype, ld.rows, ld.cols, ld.x, ld.out = L.BFLOAT16, 8192, 32000, x.data_ptr(), o.data_ptr()
        lds.append(ld)
    res["F_logsumexp_bf16_8192x32000"] = t(
        [(lambda ld=ld: L.check(lib.b200phi_logsumexp(C.byref(ld), stream()))) for ld in lds], 524304384)
    assert torch.allclose(o.float(), torch.logsumexp(xs[(21 - 1) % ROT].float(), 1), rtol=1e-2)

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

    x = xs[0]
    mx = x.max(1, keepdim=True).values
    g = torch.randn(8192, 1, device="cuda").bfloat16()
    dx = torch.empty_like(x)
    md = map_desc("equal_mul", [x, mx, g], dx)
    res["F_max_grad_bf16_8192x32000"] = t([lambda: L.check(lib.b200phi_map(C.byref(md), stream()))], 1048576000)
    assert torch.equal(dx, (x == mx).to(x.dtype) * g)
    del x, xs, dx
    torch.cuda.empty_cache()
    x = torch.rand(256, 64, 112, 112, device="cuda")
    y = torch.rand(64, 1, 1, device="cuda")
    b = torch.empty(x.shape, device="cuda", dtype=torch.bool)
    md2 = map_desc("less_than", [x, y], b)
    res["F_less_than_fp32_256x64x112x112_vs_64x1x1"] = t([lambda: L.check(lib.b200phi_map(C.byref(md2), stream()))], 1027604480)
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
    numa_node = bind_to_gpu_numa(torch, local_rank) if multi else None
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
    ws_bytes = lib.b200phi_sharded_reduce_workspace(C.byref(d))
    ws = torch.empty(max(ws_bytes, 256), dtype=torch.uint8, device="cuda")
    d.workspace, d.workspace_bytes = ws.data_ptr(), ws.numel()
    comm = reducer.comm if reducer is not None else None

    # N > 1, two routes for the [4096] exchange, both one C-ABI call per step on the reduce's stream:
    #   nccl  ncclAllReduce after the local reduce (b200phi_sharded_reduce)
    #   p2p   a 4-CTA exchange kernel over NVLink peer memory behind the unchanged reduce kernel,
    #         launched with programmatic dependent launch (b200phi_sharded_reduce_p2p)
    # B200PHI_COLLECTIVE picks one; the default is the one that measured faster (profiles/r02_multigpu.md)
    p2p = None
    collective = "none (single GPU)"
    launches_per_step = 1
    if multi:
        collective = "ncclAllReduce after the local reduce (b200phi_sharded_reduce)"
        if os.environ.get("B200PHI_COLLECTIVE", DEFAULT_COLLECTIVE) == "p2p":
            try:
                from paddle2_b200.dist import P2PReducer
                p2p = P2PReducer(max_out_numel=COLS, max_tiles=64)
                collective = ("exchange kernel over NVLink peer memory behind the reduce, programmatic "
                              "dependent launch (b200phi_sharded_reduce_p2p)")
            except Exception as exc:  # e.g. CUDA IPC not permitted in this container
                p2p = None
                collective += f" [p2p unavailable: {str(exc)[:80]}]"
        ok = torch.tensor([1 if p2p is not None else 0], device="cuda")
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)
        if int(ok.item()) == 0:
            if p2p is not None:
                p2p.close()
            p2p = None
        if p2p is not None:
            launches_per_step = 2

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

    # ---- correctness of what was timed, on EVERY rank: `out` is the last timed step's result.
    #      (1) against the closed-form global column sum, rel 1e-5; (2) all ranks hold rank 0's bits.
    want = expected_colsum(world * ROWS)
    torch.cuda.synchronize()
    if p2p is not None:
        p2p.status()
    rel = float(((out.double() - want).abs() / want).max().item())
    flags = torch.tensor([1 if rel <= 1e-5 else 0, 1], device="cuda", dtype=torch.int32)
    if multi:
        gathered = [torch.empty_like(out) for _ in range(world)]
        dist.all_gather(gathered, out)
        flags[1] = 1 if all(torch.equal(g.view(torch.int32), gathered[0].view(torch.int32)) for g in gathered) else 0
        dist.all_reduce(flags, op=dist.ReduceOp.MIN)
    result_check = {"allreduced" if multi else "local": "pass" if int(flags[0].item()) == 1 else "FAIL",
                    "ranks_identical": bool(int(flags[1].item()) == 1), "max_rel_err_rank0": rel,
                    "expected_column_sum": want,
                    "how": "closed-form inputs keyed by the global row; every rank compares its result with "
                           "the exact column sum (rel 1e-5) and with rank 0's bits"}
    if result_check.get("allreduced", result_check.get("local")) != "pass" or not result_check["ranks_identical"]:
        raise SystemExit(f"bench result check failed on the timed output: {result_check}")
    if p2p is not None:  # the peer exchange must agree with the NCCL route on the same shards
        ref = reducer.sum(x)
        torch.cuda.synchronize()
        assert torch.allclose(out, ref, rtol=1e-6), "p2p exchange result differs from the NCCL route"

    # ---- end to end through the host-buffer entry point ------------------------------------
    e2e = None
    try:
        import psutil
        avail = psutil.virtual_memory().available
    except Exception:
        avail = 0
    shard_bytes = ROWS * COLS * 4
    # a full-size pinned shard per rank when host RAM allows (NUMA-local: the process was bound to the
    # GPU's node before this allocation); else a 1 GiB staging slab streamed 8 times per step
    host_rows = ROWS if avail > (2 * world + 2) * shard_bytes else 65536
    xh = torch.empty(host_rows, COLS, dtype=torch.float32, pin_memory=True)
    for r0 in range(0, host_rows, 65536):
        xh[r0:r0 + 65536].copy_(x[r0:r0 + 65536])
    torch.cuda.synchronize()
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
    te_local = (time.perf_counter() - t0) / e2e_steps
    tt = torch.tensor([te_local], device="cuda", dtype=torch.float64)
    if multi:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    te = float(tt.item())
    # the last host pass reduced `host_rows` closed-form rows on every rank
    want_e = expected_colsum(host_rows) * world
    rel_e = float(((out_h.double() - want_e).abs() / want_e).max().item())
    assert rel_e <= 1e-5, f"e2e result check failed: rel {rel_e}"
    per_gpu_h2d = None
    if multi:
        rates = [torch.zeros(1, device="cuda", dtype=torch.float64) for _ in range(world)]
        dist.all_gather(rates, torch.tensor([shard_bytes / te_local / 1e9], device="cuda", dtype=torch.float64))
        per_gpu_h2d = [round(float(r.item()), 1) for r in rates]
    else:
        per_gpu_h2d = [round(shard_bytes / te / 1e9, 1)]
    e2e = {"value": ALGO_BYTES * world / te / 1e9, "unit": UNIT, "h2d_bytes_per_step": int(h2d),
           "d2h_bytes_per_step": int(d2h), "steps": e2e_steps,
           "per_gpu_h2d_gbs": per_gpu_h2d, "numa_node_rank0": numa_node,
           "host_buffer": f"[{host_rows},{COLS}] pinned fp32 per rank, streamed {reps}x per step in 16384-row chunks"}
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
                "traffic_source": "recorded constant: dram__bytes_read.sum + dram__bytes_write.sum of one "
                                  "ncu --set full capture of this kernel (profiles/traffic.json), not measured in this run",
                "algorithmic_bytes_per_launch": ALGO_BYTES,
                "note": "CUDA events over the timed region on the launching stream; at N>1 the step also "
                        "holds the 16 KiB exchange"}
    cpu_baseline = None
    others = None
    others_cpu = None
    if not multi:
        t_cpu, kind = cpu_sum_axis0(1, cpu_slab(16384))
        per_row = t_cpu / 16384
        rows_s = int(min(ROWS // 8, max(16384, 12.0 / per_row)))
        rows_s -= rows_s % 1024
        t_cpu, kind = cpu_sum_axis0(1, cpu_slab(rows_s))
        sb = rows_s * COLS * 4 + COLS * 4
        cpu_baseline = {"value": sb / t_cpu / 1e9, "unit": UNIT, "cores": 1, "kind": kind,
                        "sample": f"rows [0,{rows_s}) of the same shard ([{rows_s},{COLS}] fp32, {sb / 2**20:.0f} MiB), "
                                  f"one phi::funcs::SumFunctor call on Eigen::DefaultDevice, {t_cpu:.2f} s"}
        if not args.no_others:
            del x
            torch.cuda.empty_cache()
            others = other_configs(torch, pb)
            try:
                others_cpu = cpu_baselines_c1_c4()
            except Exception as exc:
                others_cpu = {"error": str(exc)[:200]}
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic (closed form, keyed by the global row)",
        "config": {"workload": WORKLOAD,
                   "l2": "inputs (8 GiB per GPU) are far larger than the 126 MB L2; no flush needed",
                   "parallelism": f"rows sharded over {world} GPU(s), one process per GPU, local reduce + "
                                  f"exchange of the [4096] partial",
                   "collective": collective},
        "clocks": sampler.summary() if sampler else None,
        "e2e": e2e, "gpu_launches": args.steps * launches_per_step,
        "roofline": roofline, "cpu_baseline": cpu_baseline, "result_check": result_check,
    }
    if others:
        line["other_configs_gbs"] = others
        line["other_configs_note"] = (f"kernel-only, CUDA-graph replay of 20-21 launches rotating over "
                                      f"{ROT} input buffers (2 for the 0.8-2 GiB inputs of C2 / C3a)")
    if others_cpu:
        line["other_configs_cpu_baseline"] = others_cpu
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
