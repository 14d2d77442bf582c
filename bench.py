#!/usr/bin/env python3
"""bench.py -- headline benchmark of the B200-native CSR SpGEMM / SpMV path.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]
                  [--workload c3|c1|c4|rmat] [--density D] [--alg 1|2|3]

One JSON line on stdout (rank 0).  A "step" is one full SpGEMM C = A @ B (symbolic +
numeric, device-resident inputs) on the workload; metric = CSR SpGEMM GFLOP/s with
FLOPs = 2 * intermediate products (what cusparseSpGEMM_getNumProducts counts,
reference cupy_cusparse/spgemm_from_txt_alg3.cu:190-191).  Default workload = BASELINE.json
configs[2], the SpGEMM_alg_comparison shape: 16384 x 16384 A @ A, uniform random, fp32, at the
heavy end of the sweep (density 1e-2).

N = 1: `value` = the device-resident step; `e2e` = the same product through the C-ABI host-buffer
call b200sp_spgemm_host (H2D of A, kernels, D2H of C inside the timed region).
N > 1 (torchrun, one rank per GPU): weak scaling - A is N stacked panels of the workload's shape,
B lives on rank 0.  `value`: every step broadcasts B over NCCL/NVLink and multiplies the rank's
panel (config.value_bcast_outside = the same with the broadcast done once during set-up).
`e2e`: the public multi-GPU entry multi.spgemm_row_panels - B broadcast, block-cyclic row panels,
C landed by every rank in ONE shared pinned host buffer at its scan offset - timed by wall clock.
extra.c5_strong (every N): the north_star multi-GPU job, R-MAT A @ A (scale 20), strong scaling,
B broadcast inside the step, partition-independent checksums against tests/golden.
--impl reference times the reference's own CPU path (scipy.sparse A @ B,
SpGEMM_vs_SpMV/profiler.py:408) on a bounded row-panel sample of the same workload.
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


M64 = (1 << 64) - 1


def build_info():
    """Version + git revision of the library build (written by __graft_entry__.build())."""
    try:
        with open(os.path.join(ROOT, "spmm_b200", "build_info.json")) as f:
            return json.load(f)
    except Exception:
        return {"git": None}


def golden_rmat(scale):
    """Partition-independent checksums of R-MAT A @ A from the CPU (tests/golden/make_rmat_checksums.py)."""
    try:
        with open(os.path.join(ROOT, "tests", "golden", "rmat_checksums.json")) as f:
            return json.load(f).get(f"rmat{scale}")
    except Exception:
        return None


def device_checksums(torch, r0, c):
    """(nnz, rowsum, colsum, colval) mod 2^64 of a device C sub-panel whose first row is r0:
    rowsum = sum (i+1) nnz(C_i), colsum = sum (col+1), colval = sum (col+1)(bits+1) with bits =
    the fp32 pattern as uint32.  Sums of sub-panels add up to the checksum of the whole product
    whatever the partition (int64 wraps = arithmetic mod 2^64)."""
    cnt = (c.indptr[1:] - c.indptr[:-1]).to(torch.int64)
    rows = torch.arange(r0 + 1, r0 + 1 + cnt.numel(), dtype=torch.int64, device=cnt.device)
    out = torch.zeros(4, dtype=torch.int64, device=cnt.device)
    out[0] = c.nnz
    out[1] = (rows * cnt).sum()
    step = 1 << 28
    for s in range(0, c.nnz, step):
        col1 = c.indices[s:s + step].to(torch.int64) + 1
        bits1 = (c.data[s:s + step].view(torch.int32).to(torch.int64) & 0xFFFFFFFF) + 1
        out[2] += col1.sum()
        out[3] += (col1 * bits1).sum()
    return out


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

def _env():
    return (int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0")))


def _max_over_ranks(torch, dist, dev, world, value):
    t = torch.tensor([float(value)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def _pin(torch, a):
    return torch.from_numpy(np.ascontiguousarray(a)).pin_memory().numpy()


def run_ours(args):
    import torch
    import torch.distributed as dist

    import spmm_b200
    from spmm_b200 import _lib, host, multi

    world, rank, local = _env()
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    B, desc = make_workload(args)   # B of the job (and the whole of A at N = 1)
    m, n = B.shape
    s = 4

    def panel_of(r):
        """Row panel r of A.  N = 1: A = B (the reference computes A @ A).  N > 1 (weak scaling):
        A = N stacked panels of the workload's shape, panel r drawn with seed 42 + r (panel 0 == B)."""
        if r == 0 or args.workload != "c3":
            return B
        return profiler_matrix(m, n, args.density or 1e-2, seed=42 + r)

    A = panel_of(rank)
    my_products = count_products(A, B)
    tp = torch.tensor([my_products], dtype=torch.int64, device=dev)
    if world > 1:
        dist.all_reduce(tp)
    total_products = int(tp.item())

    a_panel = spmm_b200.csr_matrix(A)
    a_panel._canonical = True   # preset: the reference pays its has_canonical_format kernel once per matrix, outside its timed closure too
    src = None
    if world == 1:
        b_res = a_panel
    else:
        if rank == 0:
            src = tuple(torch.from_numpy(np.ascontiguousarray(x)).to(dev) for x in multi.cast_csr_host(B))
        bp, bj, bx = multi.broadcast_csr(src, src=0, device=dev)  # also warms the communicator up
        b_meta = (bp.numel() - 1, bj.numel(), 1 if bx.dtype == torch.float64 else 0)  # known to every rank from here on
        b_res = spmm_b200.csr_matrix((bx, bj, bp), shape=(n, n))
        b_res._canonical = True
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2

    def step(bcast):
        if bcast and world > 1:
            p_, j_, x_ = multi.broadcast_csr(src, src=0, device=dev, meta=b_meta)
            b = spmm_b200.csr_matrix((x_, j_, p_), shape=(n, n))
            b._canonical = True
        else:
            b = b_res
        return spmm_b200.cusparse.spgemm(a_panel, b, alg=args.alg, chunk_fraction=args.chunk_fraction)

    def timed_steps(bcast, with_clocks):
        for _ in range(max(args.warmup, 3)):
            c = step(bcast)
        nnz_panel = c.nnz
        # bytes actually allocated for C: the single-pass mode allocates indices / data at the bound
        # sum_i min(products_i, n) and returns views of nnz(C) entries
        c_alloc = int(c.indptr.untyped_storage().nbytes() + c.indices.untyped_storage().nbytes() + c.data.untyped_storage().nbytes())
        single_pass = c.indices.untyped_storage().nbytes() > c.indices.numel() * 4
        direct = single_pass and c.indices.untyped_storage().nbytes() == c.shape[0] * c.shape[1] * 4
        del c
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        sampler = ClockSampler(local) if with_clocks else None
        if sampler:
            sampler.start()
        st0 = _lib.stats(local)
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
        ms_num, ms_sym = [], []
        t0 = time.perf_counter()
        for i in range(args.steps):
            flush.fill_(i & 0xff)  # L2 flush between timed iterations (outside the event pair)
            ev[i][0].record()
            c = step(bcast)
            ev[i][1].record()
            st = _lib.stats(local)  # waits for the numeric events of this step
            ms_num.append(st.ms_numeric)
            ms_sym.append(st.ms_symbolic)
            del c
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        wall = time.perf_counter() - t0
        clocks = sampler.stop() if sampler else None
        st1 = _lib.stats(local)
        ms = float(np.mean([a.elapsed_time(b) for a, b in ev]))
        return {"ms_job": _max_over_ranks(torch, dist, dev, world, ms), "ms_num": float(np.mean(ms_num)),
                "ms_sym": float(np.mean(ms_sym)), "wall": wall, "clocks": clocks, "nnz_panel": nnz_panel,
                "launches": int(st1.kernel_launches - st0.kernel_launches), "peak_ws": int(st1.peak_workspace_bytes),
                "c_alloc": c_alloc, "single_pass": bool(single_pass), "direct": bool(direct)}

    main_run = timed_steps(True, True)           # N > 1: B broadcast inside every step
    outside = timed_steps(False, False) if world > 1 else None
    ms_job = main_run["ms_job"]
    nnzC_panel = main_run["nnz_panel"]
    nnzC_t = torch.tensor([nnzC_panel], dtype=torch.int64, device=dev)
    if world > 1:
        dist.all_reduce(nnzC_t)
    nnzC = int(nnzC_t.item())
    value = 2.0 * total_products / (ms_job * 1e-3) / 1e9
    ms_bcast = None
    if world > 1:  # the broadcast alone
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        dist.barrier()
        e0.record()
        multi.broadcast_csr(src, src=0, device=dev, meta=b_meta)
        e1.record()
        torch.cuda.synchronize()
        ms_bcast = _max_over_ranks(torch, dist, dev, world, e0.elapsed_time(e1))

    # ---- end to end with HOST buffers ---------------------------------------------------
    e2e_steps = max(1, min(args.steps, 5))
    torch.cuda.empty_cache()
    if world == 1:
        hA = tuple(_pin(torch, x) for x in (A.indptr, A.indices, A.data))

        def e2e_step():
            cp_, cj_, cx_, _ = host.spgemm_host(hA, hA, (m, n, n), alg=args.alg, chunk_fraction=args.chunk_fraction,
                                                device=local, copy=False)
            return cp_, cj_, cx_
        e2e_api = "b200sp_spgemm_host (C ABI, pinned host CSR in, host CSR out)"
        h2d = sum(x.nbytes for x in hA)
    else:
        big = sp.vstack([panel_of(r) for r in range(world)], format="csr")
        big.sort_indices()
        hA = tuple(_pin(torch, x) for x in (big.indptr.astype(np.int32), big.indices.astype(np.int32), big.data))
        hB = multi.cast_csr_host(B) if rank == 0 else None
        big_shape = (big.shape[0], n, n)
        del big

        def e2e_step():
            return multi.spgemm_row_panels(hA, hB, big_shape, alg=args.alg, chunk_fraction=args.chunk_fraction)
        e2e_api = ("multi.spgemm_row_panels: B uploaded on rank 0 + NCCL broadcast, block-cyclic row panels H2D per rank, "
                   "C of all ranks D2H into one shared pinned host buffer at scan offsets")
        h2d = sum(x.nbytes for x in hA) + B.indptr.nbytes + B.indices.nbytes + B.data.nbytes
    e2e_step()
    e2e_step()
    te = []
    for _ in range(e2e_steps):
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        t0 = time.perf_counter()
        out = e2e_step()
        if world > 1:
            dist.barrier()
        te.append(time.perf_counter() - t0)
    t_e2e = _max_over_ranks(torch, dist, dev, world, float(np.mean(te)))
    e2e_val = 2.0 * total_products / t_e2e / 1e9
    d2h = 0
    if rank == 0:
        d2h = sum(int(x.nbytes) for x in out)
        assert int(out[0][-1]) == nnzC, f"e2e nnz(C) {int(out[0][-1])} != device-resident {nnzC}"
    del out

    extra = {}
    if not args.no_extra:
        if world > 1:
            extra["d2h_probe"] = d2h_probe(torch, dist, dev, world, rank, multi)
            extra["spmv_c2_multi"] = spmv_multi(torch, dist, dev, world, rank, spmm_b200, multi, flush)
        if not args.no_c5:
            if args.c5_ab:  # A/B inside one run: the same job without one of its two scheduling measures
                os.environ["B200SP_NO_HUB_SPLIT"] = "1"
                extra["c5_strong_no_hub_split"] = c5_strong(args, torch, dist, dev, world, rank, spmm_b200, multi, _lib, args.c5_scale)
                os.environ.pop("B200SP_NO_HUB_SPLIT", None)
                if world > 1 and not args.c5_dynamic:
                    args.c5_dynamic = True
                    extra["c5_strong_dynamic_queue"] = c5_strong(args, torch, dist, dev, world, rank, spmm_b200, multi, _lib, args.c5_scale)
                    args.c5_dynamic = False
            extra["c5_strong"] = c5_strong(args, torch, dist, dev, world, rank, spmm_b200, multi, _lib, args.c5_scale)
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    # ---- roofline of the dominant phase (numeric pass) + CPU baseline (rank 0, N = 1) ----
    peak, peak_src = hbm_peak()
    nnzA_p = int(A.nnz)
    num_bytes = spgemm_bytes(A.shape[0], nnzA_p, my_products, nnzC_panel, s)
    ms_n = main_run["ms_num"]
    achieved = num_bytes / (ms_n * 1e-3) / 1e9 if ms_n > 0 else None
    traffic, traffic_src = None, None
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            tj = json.load(f)
        traffic = tj.get(f"{args.workload}:{args.density or ''}:numeric")
        traffic_src = tj.get("_source")
    except Exception:
        pass
    out_bytes = int(4 * (A.shape[0] + 1) + nnzC_panel * (4 + s))
    in_bytes = int(4 * (A.shape[0] + 1) + A.nnz * (4 + s))
    line = {
        "metric": "csr_spgemm_gflops", "value": value, "unit": "GFLOP/s", "n_gpus": world, "steps": args.steps,
        "warmup": max(args.warmup, 3), "ms_per_step": ms_job, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": desc + (f"; weak scaling: A = {world} stacked row panels of that shape (seeds 42..{41 + world}), one per GPU, B on rank 0" if world > 1 else ""),
                   "alg": args.alg, "products": total_products, "nnzA_panel": int(A.nnz), "nnzC": nnzC,
                   "parallelism": f"row-panels x{world}" + (", B broadcast over NCCL inside every timed step (one packed collective; its shape is known to every rank from the set-up broadcast)" if world > 1 else ""),
                   "ms_broadcast_B": ms_bcast,
                   "value_bcast_outside": (2.0 * total_products / (outside["ms_job"] * 1e-3) / 1e9) if outside else None,
                   "ms_per_step_bcast_outside": outside["ms_job"] if outside else None,
                   "l2": "256 MiB buffer written between timed iterations (flush); inputs 21 MB < L2 by nature of the workload",
                   "canonical_check": "preset (_canonical = True): the has_canonical_format kernel the reference runs once per matrix is outside the timed step; B's canonical form IS verified inside every step (k_csr_check_canonical)",
                   "op_bytes_algorithmic_rank0_panel": num_bytes,
                   "ms_symbolic": main_run["ms_sym"], "ms_numeric": ms_n, "wall_s_timed_region": main_run["wall"],
                   "mode": ("single pass, direct (b200sp_spgemm_single_pass_direct: C allocated at m * n entries, one host sync)" if main_run["direct"]
                            else "single pass (b200sp_spgemm_estimate + b200sp_spgemm_single_pass)" if main_run["single_pass"] else "two pass (symbolic + numeric)"),
                   "build": build_info()},
        "clocks": main_run["clocks"],
        "e2e": {"value": e2e_val, "unit": "GFLOP/s", "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                "ms_per_step": t_e2e * 1e3, "steps": e2e_steps, "api": e2e_api},
        "gpu_launches": main_run["launches"],
        "roofline": {"bound": "hbm", "kernel": ("single-pass kernel k_num_dense_cta<T, single> (structure + values in one walk of the products; "
                                                "ms_symbolic = the estimate in front of it)" if main_run["single_pass"]
                                                else "numeric pass (all numeric-bin kernels of one step)"), "achieved": achieved,
                     "peak": peak, "peak_source": peak_src, "unit": "GB/s", "frac": (achieved / peak) if achieved else None,
                     "frac_whole_op": num_bytes / (ms_job * 1e-3) / 1e9 / peak,
                     "frac_dram": (traffic / (ms_n * 1e-3) / 1e9 / peak) if (traffic and ms_n > 0) else None,
                     "traffic": traffic, "traffic_source": traffic_src, "algorithmic_bytes_per_launch": num_bytes,
                     "note": "frac = algorithmic bytes (SURVEY 8d: every gathered B row counted once per use) / numeric-pass time; "
                             "frac_whole_op divides by the whole step (symbolic + numeric" + (" + B broadcast" if world > 1 else "") + "); "
                             "frac_dram = ncu dram bytes of the numeric kernel / its time: B is L2-resident, so DRAM sees little but C"},
        "peak_vram": {"workspace_bytes_peak": main_run["peak_ws"], "inputs_bytes": in_bytes, "output_bytes": out_bytes,
                      "output_bytes_allocated": main_run["c_alloc"],
                      "total_beyond_inputs": main_run["peak_ws"] + main_run["c_alloc"],
                      "note": "total_beyond_inputs = library workspace high-water mark + the bytes ALLOCATED for C (single-pass mode: C.indices / C.data sized at the bound sum_i min(products_i, n), of which output_bytes are used): the quantity the reference samples as dPeak VRAM with inputs resident (dense_vs_sparseGEMM/utils.py:247-250); cuSPARSE's are in extra.cusparse"},
    }
    if world == 1 and not args.no_cpu:
        rows = pick_sample_rows(A)
        P, prods = cpu_sample(A, rows)
        time_scipy(P, A, 1)
        t, _ = time_scipy(P, A, 5)
        line["cpu_baseline"] = {"value": 2.0 * prods / t / 1e9, "unit": "GFLOP/s", "cores": 1, "kind": "reference",
                                "sample": f"scipy.sparse A[:{rows}] @ A, {prods} products, median of 5, host cores available={os.cpu_count()}"}
    if world == 1 and not args.no_extra:
        extra.update(extras(args, spmm_b200, torch, dev, peak))
    if extra:
        line["extra"] = extra
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


# --------------------------------------------- multi-GPU side numbers (every rank calls) ----

def d2h_probe(torch, dist, dev, world, rank, multi):
    """Bare device-to-host bandwidth into the shared pinned buffer: every rank copies its own
    slice (<= 1 GiB) three times, all together, then rank 0 alone.  This is the ceiling `e2e`
    lives under at N GPUs (one host, one shared destination)."""
    import ctypes
    from spmm_b200 import _lib
    arena = multi._ARENA
    per = int(min(1 << 30, max(arena.size // world, 0)) // 4096 * 4096)
    if per < (1 << 20) or not arena.registered:
        return {"skipped": "no registered shared buffer"}
    arena.register_range(rank * per, per)
    if arena.unpinned:
        return {"skipped": "cudaHostRegister failed on this box"}
    buf = torch.empty(per, dtype=torch.uint8, device=dev)
    buf.fill_(rank + 1)
    lib, h = _lib.load(), _lib.handle(dev.index)
    dst = arena.base.ctypes.data + rank * per
    stream = torch.cuda.current_stream(dev)

    def copy(reps):
        for _ in range(reps):
            _lib.check(h, lib.b200sp_memcpy_async(h, ctypes.c_void_p(dst), ctypes.c_void_p(buf.data_ptr()), per, 1,
                                                  ctypes.c_void_p(stream.cuda_stream)))
        torch.cuda.synchronize()

    copy(1)
    dist.barrier()
    t0 = time.perf_counter()
    copy(3)
    t_all = _max_over_ranks(torch, dist, dev, world, time.perf_counter() - t0)
    dist.barrier()
    t_one = 0.0
    if rank == 0:
        t0 = time.perf_counter()
        copy(3)
        t_one = time.perf_counter() - t0
    dist.barrier()
    t_one = _max_over_ranks(torch, dist, dev, world, t_one)
    return {"bytes_per_rank_per_copy": per, "GBps_all_ranks_together": 3 * per * world / t_all / 1e9,
            "GBps_rank0_alone": 3 * per / t_one / 1e9,
            "note": "cudaMemcpyAsync D2H into one /dev/shm buffer page-locked by every rank (the destination of e2e)"}


def spmv_multi(torch, dist, dev, world, rank, spmm_b200, multi, flush):
    """C2 SpMV over N GPUs: row panels of A balanced by nonzeros and resident,
    every step = NCCL broadcast of x from rank 0 + the local SpMV + all-gather of the y panels."""
    M = spmv_matrix()
    m, n = M.shape
    bounds = multi.balanced_row_panels(np.diff(M.indptr).astype(np.int64), world)
    r0, r1 = bounds[rank], bounds[rank + 1]
    pp, pj, px = multi.slice_rows(M.indptr, M.indices, M.data, r0, r1)
    a = spmm_b200.csr_matrix((np.ascontiguousarray(px), np.ascontiguousarray(pj), pp), shape=(r1 - r0, n))
    a._canonical = True
    x_host = np.random.default_rng(44).standard_normal(n).astype(np.float32)
    xd = torch.from_numpy(x_host).to(dev) if rank == 0 else torch.empty(n, dtype=torch.float32, device=dev)
    width = max(bounds[r + 1] - bounds[r] for r in range(world))
    ypad = torch.zeros(width, dtype=torch.float32, device=dev)
    yall = torch.empty(width * world, dtype=torch.float32, device=dev)

    def step():
        dist.broadcast(xd, src=0)
        spmm_b200.cusparse.spmv(a, xd, y=ypad[:r1 - r0])
        dist.all_gather_into_tensor(yall, ypad)

    for _ in range(3):
        step()
    torch.cuda.synchronize()
    dist.barrier()
    ts = []
    for i in range(10):
        flush.fill_(i)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        step()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ms = _max_over_ranks(torch, dist, dev, world, float(np.median(ts)))
    y = torch.cat([yall[r * width:r * width + bounds[r + 1] - bounds[r]] for r in range(world)]).cpu().numpy()
    want = M @ x_host
    err = float((np.abs(y.astype(np.float64) - want) / ((abs(M) @ np.abs(x_host)).astype(np.float64) + 1e-30)).max())
    nbytes = 4 * (m + 1) + M.nnz * 8 + n * 4 + m * 4
    return {"ms": ms, "GBps": nbytes / ms / 1e6, "frac_hbm_of_N_gpus": nbytes / ms / 1e6 / (hbm_peak()[0] * world), "nnz": int(M.nnz),
            "max_scaled_error_vs_scipy": err, "parity_ok": bool(err <= 1e-5),
            "note": "step = NCCL broadcast of x + local SpMV of the rank's nnz-balanced row panel + NCCL all-gather of y; L2 flushed between steps"}


def c5_strong(args, torch, dist, dev, world, rank, spmm_b200, multi, _lib, scale):
    """The north_star multi-GPU job: R-MAT (scale S, edge factor 16) A @ A, STRONG scaling.  The rows
    of A are dealt in product-balanced interleaved panels (multi.plan_sub_panels), B (= A) is broadcast
    from rank 0 over NCCL INSIDE the timed step, every rank streams its sub-panels (< 2^31 entries
    each) through the GPU and folds each into partition-independent checksums (C itself is 78 GB at
    scale 20; the host landing of C is what `e2e` measures on C3).  Parity: the checksums of the whole
    product against tests/golden/rmat_checksums.json (CPU: scipy / oracle) and the first rows of every
    rank's panel bitwise against scipy, outside the timed region."""
    A = rmat_matrix(scale)
    m, n = A.shape
    At = (A.indptr, A.indices, A.data)
    rp = multi.row_products(A.indptr, A.indices, A.indptr)
    total_products = int(rp.sum())
    subs_all = multi.plan_sub_panels(rp, n, world, 4 if world > 1 else 1)
    mine = subs_all[rank]
    # --c5-dynamic: the panels are not dealt in advance but drawn from a queue shared by the ranks of the node, heaviest
    # first (8 per rank on average).  Measured in one run on 8 B200 (profiles/r02_bench_n8_v23_queue_ab.json): the rank
    # times close up (92 - 105 ms against 68 - 92 ms) but around a higher mean (98 against 79 ms) - 64 panels cost more
    # than 32 - so the job takes 105 ms instead of 92: the fixed block-cyclic deal stays the default.  (2 GPUs: 292 ms
    # against 313.)
    dynamic = world > 1 and args.c5_dynamic
    queue, panels = None, None
    if dynamic:
        panels = multi.work_ordered_panels(rp, n, 8 * world)
        queue = multi.PanelQueue()
    drawn = []
    src = None
    if world > 1 and rank == 0:
        src = tuple(torch.from_numpy(x).to(dev) for x in multi.cast_csr_host(A))
    b_res = spmm_b200.csr_matrix(A) if world == 1 else None
    if b_res is not None:
        b_res._canonical = True

    def step():
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
        ev[0].record()
        if world > 1:
            bp, bj, bx = multi.broadcast_csr(src, src=0, device=dev)
            b = spmm_b200.csr_matrix((bx, bj, bp), shape=(n, n))
            b._canonical = True
        else:
            b = b_res
        ev[1].record()
        chk = torch.zeros(4, dtype=torch.int64, device=dev)
        ms_num, cev = 0.0, []
        drawn.clear()

        def my_panels():
            if not dynamic:
                yield from mine
                return
            while True:
                i = queue.next()
                if i >= len(panels):
                    return
                drawn.append(panels[i])
                yield panels[i]

        for p0, p1 in my_panels():
            pp, pj, px = multi.slice_rows(*At, p0, p1)
            a = spmm_b200.csr_matrix((np.ascontiguousarray(px), np.ascontiguousarray(pj), pp), shape=(p1 - p0, n))
            a._canonical = True
            c = spmm_b200.cusparse.spgemm(a, b, alg=args.alg, chunk_fraction=args.chunk_fraction)
            ms_num += _lib.stats(dev.index).ms_numeric
            cev.append((torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)))
            cev[-1][0].record()
            chk += device_checksums(torch, p0, c)   # the checker, timed apart and taken out of `ms`
            cev[-1][1].record()
            del c, a
        ev[2].record()
        torch.cuda.synchronize()
        ms_chk = sum(x.elapsed_time(y) for x, y in cev)
        return chk, ev[0].elapsed_time(ev[2]) - ms_chk, ev[0].elapsed_time(ev[1]), ms_num

    if dynamic:
        queue.reset()
    step()
    if world > 1:
        dist.barrier()
    st0 = _lib.stats(dev.index)
    runs = []
    for _ in range(2):
        if dynamic:
            queue.reset()  # collective, ends with a barrier
        elif world > 1:
            dist.barrier()
        runs.append(step())
    st1 = _lib.stats(dev.index)
    if dynamic:
        mine = sorted(drawn)  # what this rank processed in the last run (for the per-rank figures and the row check below)
        queue.close()
    chk = runs[-1][0].clone()
    mine_ms = float(np.mean([r[1] for r in runs]))
    tt = torch.tensor([mine_ms, float(np.mean([r[2] for r in runs])), float(np.mean([r[3] for r in runs]))], dtype=torch.float64, device=dev)
    allt = [torch.zeros(3, dtype=torch.float64, device=dev) for _ in range(world)]
    if world > 1:
        dist.all_gather(allt, tt)
        dist.all_reduce(chk)
    else:
        allt = [tt]
    per_rank = np.array([t.cpu().numpy() for t in allt])
    got = {k: int(v) & M64 for k, v in zip(("nnz", "rowsum", "colsum", "colval"), chk.tolist())}
    # first rows of this rank's first sub-panel, bitwise against scipy (the reference's CPU path)
    rows_ok = True
    if mine:
        p0 = mine[0][0]
        p1 = min(mine[0][1], p0 + 48)
        pp, pj, px = multi.slice_rows(*At, p0, p1)
        a = spmm_b200.csr_matrix((np.ascontiguousarray(px), np.ascontiguousarray(pj), pp), shape=(p1 - p0, n))
        a._canonical = True
        b = b_res if world == 1 else spmm_b200.csr_matrix(tuple(reversed(multi.broadcast_csr(src, src=0, device=dev))), shape=(n, n))
        b._canonical = True
        c = spmm_b200.cusparse.spgemm(a, b)
        W = (A[p0:p1] @ A).tocsr()
        W.sort_indices()
        rows_ok = bool(np.array_equal(c.indptr.cpu().numpy(), W.indptr) and np.array_equal(c.indices.cpu().numpy(), W.indices)
                       and np.array_equal(c.data.cpu().numpy().view(np.uint32), W.data.view(np.uint32)))
    elif world > 1:
        multi.broadcast_csr(src, src=0, device=dev)
    ok = torch.tensor([1 if rows_ok else 0], dtype=torch.int64, device=dev)
    if world > 1:
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)
    gold = golden_rmat(scale)
    ms = float(per_rank[:, 0].max())
    peak = hbm_peak()[0]
    my_products = int(sum(rp[a:b].sum() for a, b in mine))
    nnzA_p = int(sum(A.indptr[b] - A.indptr[a] for a, b in mine))
    num_bytes = 8 * (sum(b - a for a, b in mine) + 1) + nnzA_p * 16 + my_products * 8 + int(runs[-1][0][0].item()) * 8
    ms_n = float(per_rank[rank, 2])
    return {"workload": f"C5 R-MAT scale={scale} edge_factor=16 (0.57,0.19,0.19,0.05) seed=1, A@A fp32", "scaling": "strong",
            "ms": ms, "gflops": 2.0 * total_products / ms / 1e6, "products": total_products, "nnzC": got["nnz"],
            "per_rank_ms": {"min": float(per_rank[:, 0].min()), "mean": float(per_rank[:, 0].mean()), "max": ms},
            "ms_broadcast_B": float(per_rank[:, 1].max()) if world > 1 else None,
            "sub_panels_per_rank": ([len(s_) for s_ in subs_all] if not dynamic else None),
            "panel_schedule": (f"dynamic: {len(panels)} product-balanced panels drawn heaviest-first from a queue shared by the ranks (multi.PanelQueue); rank 0 processed {len(mine)}"
                               if dynamic else "static: block-cyclic deal of 4 panels per rank (multi.plan_sub_panels)"),
            "gpu_launches_rank0": int(st1.kernel_launches - st0.kernel_launches),
            "roofline_rank0_numeric": {"achieved_GBps": num_bytes / ms_n / 1e6 if ms_n > 0 else None,
                                       "frac": num_bytes / ms_n / 1e6 / peak if ms_n > 0 else None, "ms_numeric": ms_n},
            "parity": {"checksums": got, "golden": {k: gold[k] for k in got} if gold else None,
                       "checksums_match_golden": (all(got[k] == gold[k] for k in got) if gold else None),
                       "first_rows_of_every_rank_bitwise_vs_scipy": bool(ok.item())},
            "note": "step = NCCL broadcast of B + H2D of the rank's A sub-panels + symbolic + numeric of every sub-panel; the device checksums (the parity checker) run between sub-panels and their time is subtracted; 1 warm-up, mean of 2"}


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

    world, rank, local = _env()
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
                       "ms_broadcast_B": ms_bcast, "l2": "inputs + outputs of every sub-panel exceed L2 (no flush needed)", "build": build_info()},
            "clocks": clocks, "e2e": None, "gpu_launches": int(st1.kernel_launches - st0.kernel_launches),
            "roofline": {"bound": "hbm", "kernel": "numeric pass of rank 0 (all numeric-bin kernels of its sub-panels)",
                         "achieved": num_bytes / (ms_n * 1e-3) / 1e9 if ms_n > 0 else None, "peak": peak, "peak_source": peak_src, "unit": "GB/s",
                         "frac": (num_bytes / (ms_n * 1e-3) / 1e9 / peak) if ms_n > 0 else None, "traffic": None,
                         "algorithmic_bytes_per_launch": num_bytes},
        }
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()




def cusparse_table(spmm_b200, torch, _lib):
    """cuSPARSE ALG1/2/3 (the reference's own call sequence, tools/cusparse_bench.cu = the device-resident
    twin of cupy_cusparse/spgemm_from_txt_alg{1,3}.cu:171-218) next to this library on the same inputs:
    time per SpGEMM and peak device bytes beyond the inputs (workspace + C) - the two axes of
    SpGEMM_alg_comparison/profiler.py:210-230.  Comparator only: a separate executable, never loaded
    by the product."""
    import subprocess
    import tempfile
    exe = os.path.join(ROOT, "tools", "_build", "cusparse_bench")
    if not os.path.exists(exe):
        return {"unavailable": "tools/_build/cusparse_bench not built"}
    cases = {"c1": lambda: profiler_matrix(4096, 4096, 1e-3), "c3_d1e-4": lambda: profiler_matrix(16384, 16384, 1e-4),
             "c3_d1e-3": lambda: profiler_matrix(16384, 16384, 1e-3), "c3_d1e-2": lambda: profiler_matrix(16384, 16384, 1e-2),
             "c4_d1e-2": lambda: profiler_matrix(8192, 8192, 1e-2)}
    out = {}
    lib, h = _lib.load(), _lib.handle(torch.cuda.current_device())
    for name, make in cases.items():
        A = make()
        row = {}
        with tempfile.NamedTemporaryFile(suffix=".bin", delete=False) as f:
            np.array([A.shape[0], A.shape[1], A.nnz], dtype=np.int64).tofile(f)
            A.indptr.astype(np.int32).tofile(f)
            A.indices.astype(np.int32).tofile(f)
            A.data.astype(np.float32).tofile(f)
            path = f.name
        for alg in (1, 2, 3):
            try:
                r = subprocess.run([exe, path, str(alg), "0.2", "5"], capture_output=True, text=True, timeout=300)
                j = json.loads(r.stdout.strip().splitlines()[-1])
                row[f"cusparse_alg{alg}"] = ({"error": j["error"]} if "error" in j else
                                             {"ms": j["ms_median"], "peak_bytes_beyond_inputs": j["peak_bytes_excl_inputs"]})
            except Exception as e:
                row[f"cusparse_alg{alg}"] = {"error": repr(e)[:160]}
        os.unlink(path)
        a = spmm_b200.csr_matrix(A)
        a._canonical = True
        prods = count_products(A, A)
        for alg in (1, 3):  # device-resident operator (C caller-owned: nothing to chunk, see include/b200sp.h)
            torch.cuda.synchronize()
            lib.b200sp_release_workspace(h)
            lib.b200sp_reset_stats(h)
            ts = []
            for i in range(7):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                c = spmm_b200.cusparse.spgemm(a, a, alg=alg)
                e1.record()
                torch.cuda.synchronize()
                if i >= 2:
                    ts.append(e0.elapsed_time(e1))
            st = _lib.stats(torch.cuda.current_device())
            row[f"ours_alg{alg}"] = {"ms": float(np.median(ts)), "peak_bytes_beyond_inputs": int(st.peak_workspace_bytes + c.indptr.untyped_storage().nbytes() + c.indices.untyped_storage().nbytes() + c.data.untyped_storage().nbytes()),
                                    "mode": "single pass" if c.indices.untyped_storage().nbytes() > c.indices.numel() * 4 else "two pass"}
            del c
        # host entry: ALG1 holds all of C on the device, ALG3 two chunks of it (device high-water mark, inputs included)
        from spmm_b200 import host
        hA = (A.indptr, A.indices, A.data)
        for alg in (1, 3):
            lib.b200sp_release_workspace(h)
            lib.b200sp_reset_stats(h)
            host.spgemm_host(hA, hA, A.shape + (A.shape[1],), alg=alg, chunk_fraction=0.2, copy=False)
            t0 = time.perf_counter()
            host.spgemm_host(hA, hA, A.shape + (A.shape[1],), alg=alg, chunk_fraction=0.2, copy=False)
            dt = time.perf_counter() - t0
            st = _lib.stats(torch.cuda.current_device())
            in_b = 4 * (A.shape[0] + 1) + A.nnz * 8
            row[f"ours_host_alg{alg}"] = {"ms_e2e": dt * 1e3, "device_peak_bytes_beyond_inputs": int(st.peak_workspace_bytes - in_b),
                                          "chunks": int(st.num_chunks)}
        lib.b200sp_release_workspace(h)
        row["products"], row["nnzA"] = prods, int(A.nnz)
        out[name] = row
    out["note"] = ("ms: device-resident inputs, median of 5; peak_bytes_beyond_inputs = live device allocations above the inputs "
                   "(cuSPARSE: buffers 1/2/3 + C; ours: workspace + the bytes allocated for C - alg 1 = default mode, which takes the single-pass kernel with C allocated at the bound sum_i min(products_i, n) when C is well filled; alg 3 = two passes, C allocated exactly).  ours_host_alg*: b200sp_spgemm_host wall time with host "
                   "buffers; ALG3 chunks C at chunk_fraction 0.2 so the device holds two chunks instead of C")
    return out


def extras(args, spmm_b200, torch, dev, peak):
    """Side numbers next to the headline: SpMV GB/s on C2 (all four kernels), the C3 density sweep,
    R-MAT A @ A at scale 16 / 18, and cuSPARSE ALG1/2/3 time + peak VRAM beside ours."""
    from spmm_b200 import _lib
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
        xh = np.random.default_rng(44).standard_normal(M.shape[1]).astype(np.float32)
        x = torch.from_numpy(xh).to(dev)
        y = torch.empty(M.shape[0], dtype=torch.float32, device=dev)
        nbytes = 4 * (M.shape[0] + 1) + M.nnz * 8 + M.shape[1] * 4 + M.shape[0] * 4
        for kern in ("auto", "vector4", "merge", "vector+scalar", "merge+scalar"):
            os.environ["B200SP_SPMV_KERNEL"] = kern
            ms = timed(lambda: spmm_b200.cusparse.spmv(a, x, y=y), 20)
            out[f"spmv_c2_{kern}"] = {"ms": ms, "GBps": nbytes / ms / 1e6, "frac_hbm": nbytes / ms / 1e6 / peak, "nnz": int(M.nnz)}
        os.environ.pop("B200SP_SPMV_KERNEL", None)
        out["spmv_c2_note"] = ("C2 = 2^20 square, 1.1e7 uniformly random nonzeros, fp32, L2 flushed before each run; auto = the library's "
                               "choice (row-group kernel here, 4-byte loads below 16 entries per row and 128-bit above; merge-path for skewed rows); vector4 / merge = 128-bit loads; +scalar = 4-byte loads; "
                               "algorithmic bytes = 4(m+1) + 8 nnz + 4n + 4m; x gathers (one 32 B sector per nonzero) bound it, not HBM")
        t0 = time.perf_counter()
        for _ in range(3):
            M @ xh
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
    for sc in (16, 18):  # C5 recipe at the scales one call holds
        try:
            A = rmat_matrix(sc)
            a = spmm_b200.csr_matrix(A)
            a._canonical = True
            prods = count_products(A, A)
            ms = timed(lambda: spmm_b200.cusparse.spgemm(a, a), 3)
            st = _lib.stats(dev.index)
            out[f"c5_rmat{sc}"] = {"ms": ms, "gflops": 2 * prods / ms / 1e6, "products": prods, "ms_symbolic": st.ms_symbolic, "ms_numeric": st.ms_numeric}
            del a
        except Exception as e:
            out[f"c5_rmat{sc}_error"] = repr(e)
    torch.cuda.empty_cache()
    try:
        out["cusparse"] = cusparse_table(spmm_b200, torch, _lib)
    except Exception as e:
        out["cusparse_error"] = repr(e)
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
    ap.add_argument("--c5-scale", type=int, default=20, help="R-MAT scale of extra.c5_strong")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-extra", action="store_true")
    ap.add_argument("--no-c5", action="store_true")
    ap.add_argument("--c5-dynamic", action="store_true", help="extra.c5_strong draws its panels from the shared queue (multi.PanelQueue) instead of the fixed block-cyclic deal")
    ap.add_argument("--c5-ab", action="store_true", help="also run extra.c5_strong with B200SP_NO_HUB_SPLIT=1 and with --c5-dynamic (A/B in one run)")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    elif args.workload == "rmat":
        run_rmat(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
