#!/usr/bin/env python3
"""bench.py -- headline benchmark of the B200-native CSR SpGEMM / SpMV path.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]
                  [--workload c3|c1|c4|spmv|rmat] [--density D] [--alg 1|2|3]

One JSON line on stdout (rank 0).  A "step" is one full SpGEMM C = A @ A (symbolic +
numeric, device-resident inputs) on the workload; metric = CSR SpGEMM GFLOP/s with
FLOPs = 2 * intermediate products (what cusparseSpGEMM_getNumProducts counts,
reference cupy_cusparse/spgemm_from_txt_alg3.cu:190-191).  Default workload = BASELINE.json
configs[2], the SpGEMM_alg_comparison shape: 16384 x 16384 A @ A, uniform random, fp32, at the
heavy end of the sweep (density 1e-2).  `e2e` is the same metric through the C-ABI host-buffer
call (H2D of A, kernels, D2H of C inside the timed region).  --impl reference times the
reference's own CPU path (scipy.sparse A @ B, SpGEMM_vs_SpMV/profiler.py:408) on a bounded
row-panel sample of the same workload.
"""
import argparse
import json
import os
import sys
import threading
import time

import numpy as np
import scipy.sparse as sp

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)


# ------------------------------------------------------------------------- workloads ----

def profiler_matrix(m, n, density, seed=42, dtype=np.float32):
    """SpGEMM_alg_comparison/profiler.py:146-152 (default_rng(seed), standard_normal values)."""
    rng = np.random.default_rng(seed)
    M = sp.random(m, n, density=density, format="csr", dtype=dtype, random_state=rng, data_rvs=rng.standard_normal)
    M.eliminate_zeros()
    M.sort_indices()
    return M


def spmv_matrix(n=1 << 20, density=1e-5, seed=0):
    """others/spmm/test.py:6-17 recipe with numpy's RNG: k = int(d n^2) COO draws, summed."""
    rng = np.random.default_rng(seed)
    k = int(density * n * n)
    r = rng.integers(0, n, k, dtype=np.int64)
    c = rng.integers(0, n, k, dtype=np.int64)
    v = rng.random(k, dtype=np.float32)
    M = sp.coo_matrix((v, (r, c)), shape=(n, n)).tocsr()
    M.sum_duplicates()
    M.sort_indices()
    return M


def rmat_matrix(scale=18, edge_factor=16, seed=1, abcd=(0.57, 0.19, 0.19, 0.05)):
    """R-MAT: duplicates summed, uniform[0,1) fp32 values."""
    rng = np.random.default_rng(seed)
    n, ne = 1 << scale, edge_factor << scale
    a, b, c, _ = abcd
    rows = np.zeros(ne, dtype=np.int64)
    cols = np.zeros(ne, dtype=np.int64)
    for _bit in range(scale):
        u = rng.random(ne)
        rbit = u >= a + b
        cbit = ((u >= a) & (u < a + b)) | (u >= a + b + c)
        rows = (rows << 1) | rbit
        cols = (cols << 1) | cbit
    M = sp.coo_matrix((rng.random(ne, dtype=np.float32), (rows, cols)), shape=(n, n)).tocsr()
    M.sum_duplicates()
    M.sort_indices()
    return M


def make_workload(args):
    w = args.workload
    if w == "c3":
        d = args.density or 1e-2
        return profiler_matrix(16384, 16384, d), f"C3 SpGEMM_alg_comparison 16384x16384 A@A uniform density={d:g} fp32 seed=42"
    if w == "c1":
        return profiler_matrix(4096, 4096, 1e-3), "C1 4096x4096 A@A uniform density=1e-3 fp32 seed=42"
    if w == "c4":
        d = args.density or 1e-2
        return profiler_matrix(8192, 8192, d), f"C4 dense_vs_sparseGEMM 8192x8192 A@A density={d:g} fp32"
    if w == "rmat":
        return rmat_matrix(args.scale), f"C5 R-MAT scale={args.scale} edge_factor=16 A@A fp32 seed=1"
    raise SystemExit(f"unknown workload {w}")


def count_products(A, B):
    return int(np.diff(B.indptr).astype(np.int64)[A.indices].sum())


def spgemm_bytes(m, nnzA, prods, nnzC, s=4):
    """Algorithmic bytes of one SpGEMM (BASELINE.md section 3 / SURVEY.md section 8d)."""
    return 8 * (m + 1) + nnzA * (4 + s) + 8 * nnzA + prods * (4 + s) + nnzC * (4 + s)


def hbm_peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


# ---------------------------------------------------------------------------- clocks ----

class ClockSampler:
    """Samples SM clock and throttle reasons during the timed region (NVML)."""

    def __init__(self, index):
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._th = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def _run(self):
        nv = self.nv
        names = {"hw_slowdown": 0x8, "sw_thermal_slowdown": 0x20, "hw_thermal_slowdown": 0x40,
                 "sw_power_cap": 0x4, "hw_power_brake_slowdown": 0x80}
        while not self._stop.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                for k, bit in names.items():
                    if r & bit:
                        self.reasons.add(k)
            except Exception:
                pass
            time.sleep(0.002)

    def start(self):
        if self.nv:
            self._th = threading.Thread(target=self._run, daemon=True)
            self._th.start()

    def stop(self):
        if self._th:
            self._stop.set()
            self._th.join()
        med = float(np.median(self.samples)) if self.samples else None
        return {"sm_mhz": med, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": len(self.samples)}


# ---------------------------------------------------------------------- reference arm ----

def cpu_sample(A, rows):
    """The reference's CPU path on a bounded row panel: scipy `A[:rows] @ A`."""
    P = A[:rows].tocsr()
    prods = count_products(P, A)
    return P, prods


def time_scipy(P, A, reps):
    ts = []
    for _ in range(reps):
        t0 = time.perf_counter()
        C = P @ A
        ts.append(time.perf_counter() - t0)
    return float(np.median(ts)), C.nnz


def pick_sample_rows(A, target_products=6e7):
    per_row = count_products(A, A) / max(A.shape[0], 1)
    return int(min(A.shape[0], max(1, round(target_products / max(per_row, 1)))))


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    A, desc = make_workload(args)
    rows = pick_sample_rows(A)
    P, prods = cpu_sample(A, rows)
    for _ in range(max(args.warmup, 1)):
        time_scipy(P, A, 1)
    ts = []
    for _ in range(args.steps):
        t, _nnz = time_scipy(P, A, 1)
        ts.append(t)
    t = float(np.mean(ts))
    gflops = 2.0 * prods / t / 1e9
    sample = f"scipy.sparse {sp.__name__} A[:{rows}] @ A ({prods} products per step, {rows}/{A.shape[0]} rows), 1 thread (sparsetools is single-threaded)"
    line = {
        "impl": "reference", "metric": "csr_spgemm_gflops", "value": gflops, "unit": "GFLOP/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": t * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": desc, "alg": args.alg, "host_cores_available": os.cpu_count()},
        "cpu_baseline": {"value": gflops, "unit": "GFLOP/s", "cores": 1, "kind": "reference", "sample": sample},
        "e2e": {"value": gflops, "unit": "GFLOP/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))


# ---------------------------------------------------------------------------- our arm ----

def run_ours(args):
    import torch
    import torch.distributed as dist

    import spmm_b200
    from spmm_b200 import _lib, host, multi

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    B, desc = make_workload(args)   # B of the job (and the whole of A at N = 1)
    m, n = B.shape
    s = 4
    # Row panels of A.  N = 1: A = B (the reference computes A @ A).  N > 1 (weak scaling): the job is
    # A_big @ B with A_big the vertical stack of N panels of the workload's shape, panel r drawn with
    # seed 42 + r (panel 0 == B), one panel per GPU; B lives on rank 0 and is broadcast ONCE over
    # NCCL/NVLink during set-up (timed separately: config.ms_broadcast_B), then every step is the
    # local SpGEMM of the rank's panel - no data-path collective, as Gustavson rows are independent.
    if world == 1 or args.workload != "c3":
        A = B
    else:
        A = profiler_matrix(m, n, args.density or 1e-2, seed=42 + rank)
    pp, pj, px = A.indptr, A.indices, A.data
    r0, r1 = 0, A.shape[0]
    my_products = count_products(A, B)
    tp = torch.tensor([my_products], dtype=torch.int64, device=dev)
    if world > 1:
        dist.all_reduce(tp)
    total_products = int(tp.item())

    ms_bcast = None
    if world == 1:
        b_full = spmm_b200.csr_matrix(B)
        a_panel = b_full
    else:
        src = None
        if rank == 0:
            src = tuple(torch.from_numpy(np.ascontiguousarray(x)).to(dev) for x in (B.indptr, B.indices, B.data))
        multi.broadcast_csr(src, src=0, device=dev)  # warm-up of the communicator
        torch.cuda.synchronize()
        dist.barrier()
        eb0, eb1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        eb0.record()
        bp, bj, bx = multi.broadcast_csr(src, src=0, device=dev)
        eb1.record()
        torch.cuda.synchronize()
        tb = torch.tensor([eb0.elapsed_time(eb1)], dtype=torch.float64, device=dev)
        dist.all_reduce(tb, op=dist.ReduceOp.MAX)
        ms_bcast = float(tb.item())
        b_full = spmm_b200.csr_matrix((bx, bj, bp), shape=(n, n))
        a_panel = b_full if rank == 0 else spmm_b200.csr_matrix(A)
    b_full._canonical = True
    a_panel._canonical = True

    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2

    def step():
        return spmm_b200.cusparse.spgemm(a_panel, b_full, alg=args.alg, chunk_fraction=args.chunk_fraction)

    for _ in range(max(args.warmup, 3)):
        c = step()
    nnzC_panel = c.nnz
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    sampler = ClockSampler(local)
    sampler.start()
    st0 = _lib.stats(local)
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    ms_num, ms_sym = [], []
    t_wall0 = time.perf_counter()
    for i in range(args.steps):
        flush.fill_(i & 0xff)  # L2 flush between timed iterations (outside the event pair)
        ev[i][0].record()
        c = step()
        ev[i][1].record()
        st = _lib.stats(local)  # waits for the numeric events of this step
        ms_num.append(st.ms_numeric)
        ms_sym.append(st.ms_symbolic)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    t_wall = time.perf_counter() - t_wall0
    clocks = sampler.stop()
    st1 = _lib.stats(local)
    times = [a.elapsed_time(b) for a, b in ev]
    ms = float(np.mean(times))
    tmax = torch.tensor([ms], dtype=torch.float64, device=dev)
    nnzC_t = torch.tensor([nnzC_panel], dtype=torch.int64, device=dev)
    if world > 1:
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        dist.all_reduce(nnzC_t)
    ms_job = float(tmax.item())
    nnzC = int(nnzC_t.item())
    value = 2.0 * total_products / (ms_job * 1e-3) / 1e9

    # ---- end to end through the C ABI with host buffers (pinned inputs) ------------------
    def pin(a):
        return torch.from_numpy(np.ascontiguousarray(a)).pin_memory().numpy()
    hp, hj, hx = pin(pp), pin(pj), pin(px)
    if world == 1:
        hB = None
    else:
        hB = (pin(B.indptr), pin(B.indices), pin(B.data))
    hA = (hp, hj, hx)
    e2e_steps = max(1, min(args.steps, 5))
    del c
    torch.cuda.empty_cache()

    def e2e_step():
        return host.spgemm_host(hA, hA if world == 1 else hB, (r1 - r0, n, n), alg=args.alg,
                                chunk_fraction=args.chunk_fraction, device=local, copy=False)
    e2e_step()
    e2e_step()
    if world > 1:
        dist.barrier()
    te = []
    for _ in range(e2e_steps):
        t0 = time.perf_counter()
        cp_, cj_, cx_, _ = e2e_step()
        te.append(time.perf_counter() - t0)
    t_e2e = torch.tensor([float(np.mean(te))], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t_e2e, op=dist.ReduceOp.MAX)
    e2e_val = 2.0 * total_products / float(t_e2e.item()) / 1e9
    h2d = (pp.nbytes + pj.nbytes + px.nbytes) + (0 if world == 1 else B.indptr.nbytes + B.indices.nbytes + B.data.nbytes)
    d2h = cp_.nbytes + cj_.nbytes + cx_.nbytes
    if world > 1:  # bytes of the whole job, like `value`
        tb2 = torch.tensor([h2d, d2h], dtype=torch.int64, device=dev)
        dist.all_reduce(tb2)
        h2d, d2h = int(tb2[0].item()), int(tb2[1].item())

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    # ---- roofline of the dominant phase (numeric pass) + CPU baseline (rank 0, N = 1) ----
    peak, peak_src = hbm_peak()
    nnzA_p = int(pp[-1])
    num_bytes = 8 * (r1 - r0 + 1) + nnzA_p * (4 + s) + 8 * nnzA_p + my_products * (4 + s) + nnzC_panel * (4 + s)
    ms_n = float(np.mean(ms_num))
    achieved = num_bytes / (ms_n * 1e-3) / 1e9 if ms_n > 0 else None
    traffic = None
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            traffic = json.load(f).get(f"{args.workload}:{args.density or ''}:numeric")
    except Exception:
        pass
    line = {
        "metric": "csr_spgemm_gflops", "value": value, "unit": "GFLOP/s", "n_gpus": world, "steps": args.steps,
        "warmup": max(args.warmup, 3), "ms_per_step": ms_job, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": desc + (f"; weak scaling: A = {world} stacked row panels of that shape (seeds 42..{41 + world}), one per GPU" if world > 1 else ""),
                   "alg": args.alg, "products": total_products, "nnzA_panel": int(A.nnz), "nnzC": nnzC,
                   "parallelism": f"row-panels x{world}" + (", B broadcast once over NCCL during set-up" if world > 1 else ""),
                   "ms_broadcast_B": ms_bcast,
                   "l2": "256 MiB buffer written between timed iterations (flush); inputs 21 MB < L2 by nature of the workload",
                   "op_bytes_algorithmic_rank0_panel": spgemm_bytes(m, int(A.nnz), my_products, nnzC_panel),
                   "ms_symbolic": float(np.mean(ms_sym)), "ms_numeric": ms_n, "wall_s_timed_region": t_wall},
        "clocks": clocks,
        "e2e": {"value": e2e_val, "unit": "GFLOP/s", "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                "ms_per_step": float(t_e2e.item()) * 1e3, "steps": e2e_steps},
        "gpu_launches": int(st1.kernel_launches - st0.kernel_launches),
        "roofline": {"bound": "hbm", "kernel": "numeric pass (all numeric-bin kernels of one step)", "achieved": achieved,
                     "peak": peak, "peak_source": peak_src, "unit": "GB/s", "frac": (achieved / peak) if achieved else None,
                     "traffic": traffic, "algorithmic_bytes_per_launch": num_bytes},
        "peak_vram": {"workspace_bytes_peak": int(st1.peak_workspace_bytes), "inputs_bytes": int(A.indptr.nbytes + A.indices.nbytes + A.data.nbytes),
                      "output_bytes": int(4 * (m + 1) + nnzC * (4 + s))},
    }
    if world == 1 and not args.no_cpu:
        rows = pick_sample_rows(A)
        P, prods = cpu_sample(A, rows)
        time_scipy(P, A, 1)
        t, _ = time_scipy(P, A, 5)
        line["cpu_baseline"] = {"value": 2.0 * prods / t / 1e9, "unit": "GFLOP/s", "cores": 1, "kind": "reference",
                                "sample": f"scipy.sparse A[:{rows}] @ A, {prods} products, median of 5, host cores available={os.cpu_count()}"}
    if world == 1 and not args.no_extra:
        line["extra"] = extras(args, spmm_b200, torch, dev, peak)
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


# ------------------------------------------------------- C5: R-MAT, streamed row panels ----

def run_rmat(args):
    """BASELINE config C5: R-MAT (scale S, edge factor 16) A @ A, strong scaling: the rows of A are cut
    into one panel per GPU balanced by intermediate products, B (= A) is broadcast once over NCCL, and
    every rank streams its panel through the GPU in sub-panels of < 2^31 output entries
    (multi.spgemm_panels) - nnz(C) is 1.3e9 at scale 18 and 9.7e9 at scale 20.  Each sub-panel of C is
    reduced to (nnz, sum of values) on the device and dropped, so memory stays bounded."""
    import torch
    import torch.distributed as dist

    import spmm_b200
    from spmm_b200 import _lib, multi

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    A = rmat_matrix(args.scale)
    m, n = A.shape
    rp = multi.row_products(A.indptr, A.indices, A.indptr)
    # one contiguous product-balanced panel per rank leaves rank 0 (the hub rows) 1.24x behind the
    # mean at N=8; four interleaved panels per rank even the cost per product out
    my_ranges = multi.block_cyclic_panels(rp, world, 4 if world > 1 else 1)[rank]
    total_products = int(rp.sum())
    ms_bcast = None
    if world == 1:
        b_dev = spmm_b200.csr_matrix(A)
    else:
        src = tuple(torch.from_numpy(np.ascontiguousarray(x)).to(dev) for x in (A.indptr, A.indices, A.data)) if rank == 0 else None
        multi.broadcast_csr(src, src=0, device=dev)
        torch.cuda.synchronize()
        dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        bp, bj, bx = multi.broadcast_csr(src, src=0, device=dev)
        e1.record()
        torch.cuda.synchronize()
        tb = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
        dist.all_reduce(tb, op=dist.ReduceOp.MAX)
        ms_bcast = float(tb.item())
        b_dev = spmm_b200.csr_matrix((bx, bj, bp), shape=(n, n))
    b_dev._canonical = True
    At = (A.indptr, A.indices, A.data)

    def step():
        nnz, chk, npan, ms_num = 0, torch.zeros((), dtype=torch.float64, device=dev), 0, 0.0
        for r0, r1 in my_ranges:
            for _p0, _p1, c in multi.spgemm_panels(At, b_dev, (m, n, n), r0, r1, alg=args.alg, chunk_fraction=args.chunk_fraction, row_prods=rp):
                nnz += c.nnz
                chk += c.data.sum(dtype=torch.float64)
                npan += 1
                ms_num += _lib.stats(local).ms_numeric
                del c
        return nnz, chk, npan, ms_num

    for _ in range(max(1, min(args.warmup, 2))):
        step()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    sampler = ClockSampler(local)
    sampler.start()
    st0 = _lib.stats(local)
    ts, ms_nums = [], []
    for _ in range(args.steps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        nnz, chk, npan, ms_num = step()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
        ms_nums.append(ms_num)
    if world > 1:
        dist.barrier()
    clocks = sampler.stop()
    st1 = _lib.stats(local)
    t = torch.tensor([float(np.mean(ts))], dtype=torch.float64, device=dev)
    agg = torch.tensor([nnz, npan], dtype=torch.int64, device=dev)
    chk_t = chk.clone()
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(agg)
        dist.all_reduce(chk_t)
    ms_job = float(t.item())
    if rank == 0:
        peak, peak_src = hbm_peak()
        my_products = int(sum(rp[a:b].sum() for a, b in my_ranges))
        nnzA_p = int(sum(A.indptr[b] - A.indptr[a] for a, b in my_ranges))
        my_rows = int(sum(b - a for a, b in my_ranges))
        num_bytes = 8 * (my_rows + 1) + nnzA_p * 16 + my_products * 8 + nnz * 8
        ms_n = float(np.mean(ms_nums))
        line = {
            "metric": "csr_spgemm_gflops", "value": 2.0 * total_products / (ms_job * 1e-3) / 1e9, "unit": "GFLOP/s", "n_gpus": world,
            "steps": args.steps, "warmup": max(1, min(args.warmup, 2)), "ms_per_step": ms_job, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"C5 R-MAT scale={args.scale} edge_factor=16 (0.57,0.19,0.19,0.05) seed=1, A@A fp32", "alg": args.alg,
                       "rows": m, "nnzA": int(A.nnz), "products": total_products, "nnzC": int(agg[0].item()),
                       "sub_panels_all_ranks": int(agg[1].item()), "checksum_sum_of_values": float(chk_t.item()),
                       "parallelism": f"row panels x{world} balanced by products ({len(my_ranges)} interleaved panels per rank), B broadcast once over NCCL; C streamed in sub-panels of < 2^31 entries and reduced on device",
                       "ms_broadcast_B": ms_bcast, "l2": "inputs + outputs of every sub-panel exceed L2 (no flush needed)"},
            "clocks": clocks, "e2e": None, "gpu_launches": int(st1.kernel_launches - st0.kernel_launches),
            "roofline": {"bound": "hbm", "kernel": "numeric pass of rank 0 (all numeric-bin kernels of its sub-panels)",
                         "achieved": num_bytes / (ms_n * 1e-3) / 1e9 if ms_n > 0 else None, "peak": peak, "peak_source": peak_src, "unit": "GB/s",
                         "frac": (num_bytes / (ms_n * 1e-3) / 1e9 / peak) if ms_n > 0 else None, "traffic": None,
                         "algorithmic_bytes_per_launch": num_bytes},
        }
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()



def extras(args, spmm_b200, torch, dev, peak):
    """Side numbers next to the headline: SpMV GB/s on C2, the C3 density sweep and R-MAT A @ A at
    scale 16 / 18 (few reps)."""
    out = {}
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

    def timed(fn, reps):
        fn()
        fn()
        ts = []
        for i in range(reps):
            flush.fill_(i)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            fn()
            e1.record()
            torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1))
        return float(np.median(ts))

    try:
        M = spmv_matrix()
        a = spmm_b200.csr_matrix(M)
        a._canonical = True
        x = torch.from_numpy(np.random.default_rng(44).standard_normal(M.shape[1]).astype(np.float32)).to(dev)
        y = torch.empty(M.shape[0], dtype=torch.float32, device=dev)
        nbytes = 4 * (M.shape[0] + 1) + M.nnz * 8 + M.shape[1] * 4 + M.shape[0] * 4
        for kern in ("merge", "vector"):  # merge-path is the default; vector = group-of-lanes-per-row
            os.environ["B200SP_SPMV_KERNEL"] = kern
            ms = timed(lambda: spmm_b200.cusparse.spmv(a, x, y=y), 20)
            out[f"spmv_c2_{kern}"] = {"ms": ms, "GBps": nbytes / ms / 1e6, "frac_hbm": nbytes / ms / 1e6 / peak, "nnz": int(M.nnz)}
        out["spmv_c2_note"] = ("C2 = 2^20 square, 1.1e7 uniformly random nonzeros, fp32, L2 flushed before each run; bound by the "
                               "L1 tag stage (random 4-byte gathers of x: ~1.27 SM-cycles each, tools/gather_probe.cu -> ~50 us), "
                               "not by HBM; algorithmic bytes = 4(m+1) + 8 nnz + 4n + 4m")
        os.environ.pop("B200SP_SPMV_KERNEL", None)
        t0 = time.perf_counter()
        for _ in range(3):
            M @ x.cpu().numpy()
        out["spmv_c2_scipy_ms"] = (time.perf_counter() - t0) / 3 * 1e3
        del a, x, y
    except Exception as e:  # extras never sink the headline
        out["spmv_error"] = repr(e)
    for d in (1e-4, 1e-3, 1e-2):
        try:
            A = profiler_matrix(16384, 16384, d)
            a = spmm_b200.csr_matrix(A)
            a._canonical = True
            prods = count_products(A, A)
            ms = timed(lambda: spmm_b200.cusparse.spgemm(a, a), 5)
            out[f"c3_d{d:g}"] = {"ms": ms, "gflops": 2 * prods / ms / 1e6, "products": prods}
        except Exception as e:
            out[f"c3_d{d:g}_error"] = repr(e)
    for sc in (16, 18):  # C5 recipe at the scales one call holds (the scale-22 job is `--workload rmat --scale 22`)
        try:
            A = rmat_matrix(sc)
            a = spmm_b200.csr_matrix(A)
            a._canonical = True
            prods = count_products(A, A)
            ms = timed(lambda: spmm_b200.cusparse.spgemm(a, a), 3)
            out[f"c5_rmat{sc}"] = {"ms": ms, "gflops": 2 * prods / ms / 1e6, "products": prods}
            del a
        except Exception as e:
            out[f"c5_rmat{sc}_error"] = repr(e)
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c3", choices=["c3", "c1", "c4", "rmat"])
    ap.add_argument("--density", type=float, default=None)
    ap.add_argument("--scale", type=int, default=18)
    ap.add_argument("--alg", type=int, default=1)
    ap.add_argument("--chunk-fraction", type=float, default=0.2)
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-extra", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    elif args.workload == "rmat":
        run_rmat(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
