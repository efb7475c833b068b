#!/usr/bin/env python
"""bench.py -- elements/s of the (Ke + f_int) global assembly hot path on B200.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--n CUBE] [--impl reference]

Workload (BASELINE.json configs[1]): synthetic structured hex8 cube n^3 (default 100^3 =
1M elements, 3.09M dof, 245M nonzeros) per GPU, linear elastic, jittered interior nodes so
elements are not congruent.  A step = one femx_assemble_K_fint call: every CSR value of K and
the internal force f = sum_e Ke u_e are (re)computed from coordinates and u.  `value` is
device-timed with inputs resident in HBM; `e2e` goes through the same C-ABI call with pinned
HOST buffers (u H2D, K values + f D2H inside the timed region).
N > 1 (weak scaling): the (Px n) x (Py n) x (Pz n) cube in Px x Py x Pz blocks of n^3 elements, one per GPU -- 2x1x1,
2x2x1, 2x2x2; N = 8 is BASELINE configs[4] (200^3, three interface faces per rank) -- node numbering contiguous per
block, rows owned by the PETSC_DECIDE split.  The step then includes femx's interface exchange: remote stores into the
owners' arenas over NVLink, the step barrier among neighbours and the receiver add over the interface rows.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

BYTES_PER_EL = {"hex8": 2074.0}
KERNEL_NAME = {0: "hex_mma_kernel", 1: "elem_rows_kernel + gather_rows_kernel", 9: "fused_rows_kernel", 12: "hex_sym6_kernel", 20: "hex_mma_kernel"}
# conn 32 + mat 4 + coords 24.73 + u 24.73 + K 1963.5 + R 24.73


def algorithmic_bytes(m, nnz):
    """Compulsory traffic of one K + f_int pass: every input read once, every output written once."""
    return (m.eind.nbytes + m.emat.nbytes + m.coords.nbytes + m.coords.nbytes  # conn, mat, coords, u
            + 8.0 * nnz + 8.0 * m.pdim * m.nnds)                                # K values, R


class ClockSampler(threading.Thread):
    def __init__(self, index=0):
        super().__init__(daemon=True)
        self.index, self.rows, self.stop_flag = index, [], False

    def run(self):
        q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
            "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
        while not self.stop_flag:
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={q}", "--format=csv,noheader,nounits", "-i", str(self.index)],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.rows.append([x.strip() for x in out.split(",")])
            except Exception:
                pass
            time.sleep(0.1)

    def summary(self):
        if not self.rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        sm = sorted(float(r[0]) for r in self.rows if r[0].replace(".", "").isdigit())
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(r[3 + i].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": float(self.rows[0][1]), "reasons": reasons,
                "samples": len(self.rows)}


def build_workload(n, nranks, rank):
    from fem_b200 import meshgen as mg
    if nranks == 1:
        m = mg.jitter(mg.hex_cube(n, clamp=False))
    else:
        # weak scaling: one n^3 block of the (Px n) x (Py n) x (Pz n) cube per rank; each rank assembles the rows of its own
        # elements (partition by element, rows by the PETSC_DECIDE split)
        m = mg.hex_block_rank(n, mg.block_grid(nranks), rank)
    return m, mg.displacement_field(m)


def bind_to_gpu_numa_node(index):
    """Pin this rank's host threads (and therefore its pinned buffers' first touch) to the CPUs NVML names as local to the
    GPU: eight ranks copying 2 GB each through one socket is what made the round-1 e2e collapse at N = 8."""
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(index)
        words = pynvml.nvmlDeviceGetCpuAffinity(h, (os.cpu_count() + 63) // 64)
        cpus = {64 * w + b for w, word in enumerate(words) for b in range(64) if (word >> b) & 1}
        allowed = os.sched_getaffinity(0)
        if cpus & allowed:
            os.sched_setaffinity(0, cpus & allowed)
            return len(cpus & allowed)
    except Exception:
        pass
    return 0


def other_configs(args, api, mg):
    """N = 1 only: the BASELINE configurations that are not the headline, each with device ms (median of 5), elements/s,
    algorithmic GB/s and its fraction of the measured HBM rate (SURVEY.md 8(d) bytes per element)."""
    out = []
    try:
        peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))).get("hbm_gbs", 6650.0))
    except Exception:
        peak = 6650.0

    def timed(fx, call, reps=5):
        ms = []
        for _ in range(reps + 1):
            call()
            ms.append(fx.last_kernel_ms)
        return float(np.median(ms[1:]))

    def add(name, nels, ms, bpe, **kw):
        gbs = nels * bpe / (ms * 1e-3) / 1e9
        out.append(dict(config=name, ms=ms, elements_per_s=nels / (ms * 1e-3), algorithmic_bytes_per_element=bpe, algorithmic_GBps=gbs,
                        frac=gbs / peak, **kw))

    # hex8 100^3 through the other kernels: the DMMA tile kernel and the general fused rows kernel (what an off-grid hex8 mesh gets)
    m = mg.jitter(mg.hex_cube(args.n, clamp=False))
    u = mg.displacement_field(m)
    for var, nm in ((12, "hex_sym6_kernel (variant 12)"), (9, "fused_rows_kernel (variant 9: any hex8 mesh)")):
        with api.Femx(assembly_variant=var) as fx:
            fx.set_mesh(m).build_plan()
            d_u = fx.device_buffer(u.nbytes).upload(u); d_v = fx.device_buffer(8 * fx.nnz); d_f = fx.device_buffer(8 * fx.nrows)
            ms = timed(fx, lambda: fx._ck(fx.L.femx_assemble_K_fint(fx.h, d_u.ptr, d_v.ptr, d_f.ptr, api.DEVICE)))
            add(f"C2 hex8 {args.n}^3 K + f_int, {nm}", m.nels, ms, 2074.0)
            if var == 9:
                d_s = fx.device_buffer(8 * 6 * m.nlnds)
                ms = timed(fx, lambda: fx._ck(fx.L.femx_stress(fx.h, d_u.ptr, d_s.ptr, api.ROW_MAJOR, api.DEVICE)))
                add(f"C2 hex8 {args.n}^3 nodal-averaged stress", m.nels, ms, 32 + 4 + 2 * 24.73 + 48 * m.nlnds / m.nels)
    m = mg.tet_cube(94)
    u = mg.displacement_field(m)
    with api.Femx() as fx:
        fx.set_mesh(m).build_plan()
        d_u = fx.device_buffer(u.nbytes).upload(u); d_v = fx.device_buffer(8 * fx.nnz); d_f = fx.device_buffer(8 * fx.nrows)
        d_s = fx.device_buffer(8 * 6 * m.nlnds)
        ms = timed(fx, lambda: fx._ck(fx.L.femx_assemble_K_fint(fx.h, d_u.ptr, d_v.ptr, d_f.ptr, api.DEVICE)))
        add("C3 tet4 94^3x6 K + f_int (rows_simplex_kernel)", m.nels, ms, 215.0)
        ms = timed(fx, lambda: fx._ck(fx.L.femx_stress(fx.h, d_u.ptr, d_s.ptr, api.ROW_MAJOR, api.DEVICE)))
        add("C3 tet4 94^3x6 nodal-averaged stress", m.nels, ms, 36.5)
    m = mg.cohesive_strip(1000, 1000)
    nn = m.nlnds
    u = np.zeros((nn, 2)); u[nn // 2:, 1] = 1e-3
    with api.Femx(quirks=api.Q_UEL) as fx:
        fx.set_mesh(m).build_plan()
        d_u = fx.device_buffer(u.nbytes).upload(u); d_k = fx.device_buffer(8 * fx.nnz); d_j = fx.device_buffer(8 * fx.nnz)
        d_f = fx.device_buffer(8 * fx.nrows); d_r = fx.device_buffer(8 * fx.nrows)
        d_du = fx.device_buffer(8 * fx.nrows).upload(np.zeros(fx.nrows)); d_fe = fx.device_buffer(8 * fx.nrows).upload(np.zeros(fx.nrows))
        ms = timed(fx, lambda: fx._ck(fx.L.femx_assemble_K_fint(fx.h, d_u.ptr, d_k.ptr, d_f.ptr, api.DEVICE)))
        add("C4 quad4 2x1000^2 + 1000 coh K + f_int (fused_rows_kernel)", m.nels, ms, 356.0)
        ms = timed(fx, lambda: fx._ck(fx.L.femx_jacobian(fx.h, d_k.ptr, d_u.ptr, float(m.dt), d_j.ptr, api.DEVICE)))
        add("C4 femx_jacobian (copy K, cohesive overlay, BC diagonal)", m.nels, ms, 16.0 * fx.nnz / m.nels)
        ms = timed(fx, lambda: fx._ck(fx.L.femx_jacobian_update(fx.h, d_k.ptr, d_u.ptr, float(m.dt), d_j.ptr, api.DEVICE)))
        add("C4 femx_jacobian_update (cohesive blocks + BC diagonal only)", m.nels, ms, 0.0)
        ms = timed(fx, lambda: fx._ck(fx.L.femx_residual(fx.h, d_j.ptr, d_u.ptr, d_du.ptr, d_fe.ptr, float(m.dt), float(m.t), d_r.ptr, api.DEVICE)))
        add("C4 femx_residual (cohesive forces, BC rows, J du - f)", m.nels, ms, (12.0 * fx.nnz + 24.0 * fx.nrows) / m.nels)
    return out


def cpu_sample(n_cpu, threads):
    """Oracle (CPU restatement of src/main.F90:157-166 + f = Ke u_e) on a bounded hex8 sample."""
    from fem_b200 import meshgen as mg
    from oracle import pyoracle as orc
    m = mg.jitter(mg.hex_cube(n_cpu, clamp=False))
    u = mg.displacement_field(m)
    om = orc.Mesh(m)
    # the reference preallocates pdim*(nodal_bw+1) slots per row (src/main.F90:153); capped at the true hex8 row length so
    # that a 56^3 sample fits in host memory (a smaller reservation only costs the CPU arm reallocations it would not have)
    sec, nnz = orc.time_assembly(om, u.reshape(-1), prealloc=min(m.pdim * (m.nodal_bw + 1), 81), nthreads=threads)
    return m.nels, sec


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    from oracle import pyoracle as orc
    orc.build()
    threads = os.cpu_count() or 1
    try:
        threads = len(os.sched_getaffinity(0))
    except Exception:
        pass
    n_cpu = args.cpu_n
    for _ in range(args.warmup):
        cpu_sample(max(8, n_cpu // 2), threads)
    t_tot, els = 0.0, 0
    for _ in range(args.steps):
        ne, sec = cpu_sample(n_cpu, threads)
        els += ne
        t_tot += sec
    val = els / t_tot
    line = {"impl": "reference", "metric": "elements/sec (Ke+f_int assembly)", "value": val, "unit": "elements/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * t_tot / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": f"hex8 cube {args.n}^3 per GPU (K + f_int)", "sample": f"hex8 cube {n_cpu}^3 per step"},
            "cpu_baseline": {"value": val, "unit": "elements/s", "cores": threads, "kind": "port",
                             "sample": f"{args.steps} x hex8 {n_cpu}^3 ({n_cpu ** 3} elements), oracle/ CPU restatement of "
                                       "src/main.F90:157-166 (reference itself is Fortran+PETSc: not buildable here)"},
            "e2e": {"value": val, "unit": "elements/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--exchange", choices=("peer", "peer-fused", "nccl"), default="peer",
                    help="N > 1: interface rows pushed into the owner's arena over NVLink by femx's own kernel behind the assembly "
                         "kernel (peer), stored there by the assembly kernel itself (peer-fused), or sent with NCCL after it")
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--n", type=int, default=100, help="hex8 cube edge per GPU")
    ap.add_argument("--cpu-n", type=int, default=56, help="edge of the bounded CPU sample cube")
    ap.add_argument("--impl", default="femx")
    ap.add_argument("--variant", type=int, default=0)
    ap.add_argument("--lz", type=int, default=0)
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-other", action="store_true", help="skip the other_configs array (C3 / C4 / other hex8 kernels; N = 1 only)")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    numa_cpus = bind_to_gpu_numa_node(local_rank) if world > 1 else 0
    import torch
    import torch.distributed as dist
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (femx has no CPU fallback)")
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    import ctypes as C
    from fem_b200 import api, meshgen as mg
    m, u = build_workload(args.n, world, rank)
    fx = api.Femx(device=local_rank, quirks=api.Q_REFERENCE, assembly_variant=args.variant, lz=args.lz, peer_in_kernel=args.exchange == "peer-fused")
    t0 = time.time()
    fx.set_mesh(m).build_plan()
    plan_s = time.time() - t0
    nnz, nrows = fx.nnz, fx.nrows

    d_u = fx.device_buffer(u.nbytes).upload(u)
    d_f = fx.device_buffer(8 * nrows)
    da = None
    if world > 1:
        from fem_b200 import parallel as par
        da = par.DistributedAssembly(fx, dist, torch.device("cuda", local_rank), peer=args.exchange != "nccl")
        vals_ptr = C.c_void_p(da.vals_local.data_ptr())
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    else:
        d_vals = fx.device_buffer(8 * nnz)
        vals_ptr = d_vals.ptr

    def step_device():
        # peer mode: kernel + remote stores + neighbour barrier + receiver add (interface rows) are all inside this call
        fx._ck(fx.L.femx_assemble_K_fint(fx.h, d_u.ptr, vals_ptr, d_f.ptr, api.DEVICE))
        ms = fx.last_kernel_ms
        asm_ms[0] += fx.L.femx_last_assembly_kernel_ms(fx.h)
        if da is not None and not da.peer:
            ev0.record()
            da.exchange()
            ev1.record()
            ev1.synchronize()
            ms += ev0.elapsed_time(ev1)
        return ms

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    asm_ms = [0.0]                       # device time of the assembly kernel alone (the rest of a step: push, neighbours' barrier, add)
    sampler = ClockSampler(local_rank)
    sampler.start()                      # samples from warm-up to the end of the e2e loop: the GPU is under load throughout
    for _ in range(args.warmup):
        step_device()
    barrier()
    l0 = fx.launches
    asm_ms[0] = 0.0
    w0 = time.perf_counter()
    dev_ms = 0.0
    for _ in range(args.steps):
        dev_ms += step_device()
    barrier()
    wall_ms = 1e3 * (time.perf_counter() - w0)
    launches = fx.launches - l0

    # e2e: pinned host buffers through the C ABI.  N = 1: one femx_assemble_K_fint(FEMX_HOST) call per step.  N > 1: the rank's
    # OWNED part travels -- u in, then the owned value slice, the merged interface rows and the owned internal force out.
    e2e = None
    if not args.no_e2e:
        hu = fx.pinned(u.size); hu.array[:] = u.reshape(-1)
        if da is None:
            hv = fx.pinned(nnz); hf = fx.pinned(nrows)

            def step_host():
                fx._ck(fx.L.femx_assemble_K_fint(fx.h, hu.ptr, hv.ptr, hf.ptr, api.HOST))
            d2h = 8 * nnz + 8 * nrows
        else:
            ep = da.ep
            n_own, n_mer, n_f = ep.own_hi - ep.own_lo, len(ep.m_own), ep.row_end - ep.row_start
            hv = fx.pinned(n_own + n_mer); hf = fx.pinned(n_f)
            own_src = C.c_void_p(da.vals_local.data_ptr() + 8 * ep.own_lo)
            f_src = C.c_void_p(d_f.ptr.value + 8 * ep.row_start)
            hv_mer = C.c_void_p(hv.ptr.value + 8 * n_own)

            def step_host():
                fx._ck(fx.L.femx_memcpy(fx.h, d_u.ptr, hu.ptr, u.nbytes, api.DEVICE, api.HOST))
                fx._ck(fx.L.femx_assemble_K_fint(fx.h, d_u.ptr, vals_ptr, d_f.ptr, api.DEVICE))
                if not da.peer:
                    da.exchange()
                fx._ck(fx.L.femx_memcpy(fx.h, hv.ptr, own_src, 8 * n_own, api.HOST, api.DEVICE))
                if da.peer and n_mer:
                    p, nn_ = C.c_void_p(), C.c_int64()
                    fx._ck(fx.L.femx_peer_merged(fx.h, C.byref(p), C.byref(nn_)))
                    fx._ck(fx.L.femx_memcpy(fx.h, hv_mer, p, 8 * n_mer, api.HOST, api.DEVICE))
                fx._ck(fx.L.femx_memcpy(fx.h, hf.ptr, f_src, 8 * n_f, api.HOST, api.DEVICE))
            d2h = 8 * (n_own + n_mer + n_f)
        step_host()
        barrier()
        e0 = time.perf_counter()
        ke = max(2, min(args.steps, 5))
        for _ in range(ke):
            step_host()
        barrier()
        e2e_ms = 1e3 * (time.perf_counter() - e0) / ke
        # spot check of what came back (size-independent; tests/ hold the oracle parity)
        assert np.isfinite(hv.array[:1000]).all() and np.isfinite(hf.array).all()
        e2e = {"ms": e2e_ms, "h2d": int(u.nbytes), "d2h": int(d2h)}
        hu.free(); hv.free(); hf.free()
    sampler.stop_flag = True
    sampler.join(timeout=2)

    # in-bench correctness check (N = 1): the timed kernel against the general fused rows kernel, an independent algorithm
    check = None
    if world == 1 and args.variant != 9:
        a = torch.as_tensor(par_dev(vals_ptr.value, nnz), device="cuda")
        fa = torch.as_tensor(par_dev(d_f.ptr.value, nrows), device="cuda")
        with api.Femx(device=local_rank, assembly_variant=9) as f9:
            f9.set_mesh(m).build_plan()
            d_v9 = f9.device_buffer(8 * nnz); d_f9 = f9.device_buffer(8 * nrows); d_u9 = f9.device_buffer(u.nbytes).upload(u)
            f9._ck(f9.L.femx_assemble_K_fint(f9.h, d_u9.ptr, d_v9.ptr, d_f9.ptr, api.DEVICE))
            b = torch.as_tensor(par_dev(d_v9.ptr.value, nnz), device="cuda")
            fb = torch.as_tensor(par_dev(d_f9.ptr.value, nrows), device="cuda")
            check = {"K_max_diff_vs_general_kernel_rel": float(((a - b).abs().max() / b.abs().max()).item()),
                     "f_max_diff_vs_general_kernel_rel": float(((fa - fb).abs().max() / fb.abs().max()).item()),
                     "sum_K": float(a.sum().item()), "sum_f": float(fa.sum().item())}
            del b, fb
        del a, fa

    if world > 1 and da is not None and da.peer:
        # in-bench check (N > 1), size-independent: a rigid translation is in the null space of the assembled matrix, so in every
        # MERGED interface row (own part + what the peers pushed, added by femx) the entries of each column component sum to zero;
        # a lost, duplicated or misplaced contribution of the exchange leaves a residual of the size of an entry
        ep = da.ep
        mer = da.merged().cpu().numpy()
        nrow = len(ep.if_rows)
        worst = 0.0
        if len(mer):
            rows = np.repeat(np.arange(nrow, dtype=np.int64), np.diff(ep.if_rowptr))
            sums = np.bincount(rows * m.pdim + (ep.if_colidx.astype(np.int64) % m.pdim), weights=mer, minlength=nrow * m.pdim)
            worst = float(np.abs(sums).max() / np.abs(mer).max())
        tw = torch.tensor([worst, float(nrow)], device="cuda", dtype=torch.float64)
        tmx = tw.clone(); dist.all_reduce(tmx, op=dist.ReduceOp.MAX)
        tsm = tw.clone(); dist.all_reduce(tsm)
        check = {"merged_interface_rows_rigid_translation_residual_rel": float(tmx[0]), "merged_interface_rows_all_ranks": int(tsm[1]),
                 "note": "max over ranks and rows of |sum_j K[row, (node j, component c)]| / max|K|, rows merged by femx_assemble_K_fint"}
        assert float(tmx[0]) < 1e-9, check

    t = torch.tensor([dev_ms, wall_ms, e2e["ms"] if e2e else 0.0], device="cuda", dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dev_ms_rank = dev_ms
    dev_ms, wall_ms, e2e_ms = (float(x) for x in t.tolist())
    breakdown = None
    if world > 1:   # where a multi-GPU step goes: the kernel itself (per-GPU spread) vs push + waiting for the slowest neighbour + add
        k = asm_ms[0] / args.steps
        x = dev_ms_rank / args.steps - k
        tt = torch.tensor([k, -k, x, -x, k, x], device="cuda", dtype=torch.float64)
        hi = tt[:4].clone(); dist.all_reduce(hi, op=dist.ReduceOp.MAX)
        sm = tt[4:].clone(); dist.all_reduce(sm)
        breakdown = {"assembly_kernel_ms": {"min": -float(hi[1]), "mean": float(sm[0]) / world, "max": float(hi[0])},
                     "push_barrier_add_ms": {"min": -float(hi[3]), "mean": float(sm[1]) / world, "max": float(hi[2])},
                     "note": "per rank and step; assembly_kernel_ms = device time until the last tile kernel ends, push_barrier_add_ms = what is left of the "
                             "exchange after it (the exchange runs on a high-priority stream behind the interface tiles, overlapped with the interior "
                             "tiles; FEMX_PEER_OVERLAP=0: serial behind the kernel); the barrier waits for the slowest neighbour"}
    io = torch.tensor([float(e2e["h2d"]) if e2e else 0.0, float(e2e["d2h"]) if e2e else 0.0], device="cuda", dtype=torch.float64)
    if world > 1:
        dist.all_reduce(io)             # bytes of all ranks per step
    nels_total = m.nels * world          # every rank holds an n^3 block
    if rank == 0:
        ms_step = dev_ms / args.steps
        value = nels_total / (ms_step * 1e-3)
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        peak = float(peaks.get("hbm_gbs", 6650.0))
        alg = algorithmic_bytes(m, nnz)
        achieved = alg / (ms_step * 1e-3) / 1e9
        traffic, pipes = None, None
        try:   # dram__bytes_read.sum + dram__bytes_write.sum of the dominant kernel from the committed ncu capture
            tj = json.load(open(os.path.join(ROOT, "profiles", "r02_traffic.json")))
            if args.n == 100:
                ent = tj.get(KERNEL_NAME.get(args.variant, "?"), {})
                traffic = ent.get("traffic_bytes_per_launch")
                if "lsu_data_pipe_pct_of_peak" in ent:   # what actually bounds the kernel (committed ncu capture, not measured live)
                    pipes = {k: ent[k] for k in ("lsu_data_pipe_pct_of_peak", "fp64_pipe_pct_of_peak", "dram_pct_of_peak") if k in ent}
                    pipes["source"] = ent.get("source", "profiles/")
        except Exception:
            pass
        if world > 1:
            grid = mg.block_grid(world)
            par_txt = (f"{grid[0]}x{grid[1]}x{grid[2]} blocks of {args.n}^3 elements (N = 8: BASELINE configs[4]), node numbering contiguous per block, "
                       "PETSC_DECIDE row split, interface rows "
                       + ((("stored by the assembly kernel" if args.exchange == "peer-fused" else "pushed by femx's peer kernel")
                           + " into the owner's arena over NVLink + device-side step barrier among neighbours + receiver add over the "
                             "interface rows, all inside the timed step"
                           + ("" if args.exchange == "peer-fused" else "; tiles owning interface rows first, exchange overlapped with the interior tiles"))
                          if da.peer else "sent with NCCL all_to_all after the kernel (receiver add on the host, not timed)")
                       + f" ({int(da.ep.recv_counts.sum()) * 8} B received, {len(da.ep.m_own)} merged values on rank 0 per step)")
        else:
            par_txt = "1 GPU"
        line = {"metric": "elements/sec (Ke+f_int assembly)", "value": value, "unit": "elements/s", "n_gpus": world,
                "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                "config": {"workload": f"hex8 cube {args.n}^3 per GPU, jittered, K + f_int (BASELINE configs[1])",
                           "elements_per_gpu": m.nels, "dof_per_gpu": int(m.pdim * m.nlnds), "nnz_per_gpu": int(nnz),
                           "l2": "outputs (8*nnz bytes) exceed the 126 MB L2; no flush needed",
                           "parallelism": par_txt,
                           "variant": args.variant, "lz": args.lz, "plan_build_s": round(plan_s, 3), "plan_pass_s": round(fx.L.femx_plan_seconds(fx.h), 4), "wall_ms_per_step": wall_ms / args.steps},
                "gpu_launches": int(launches), "step_breakdown": breakdown,
                "clocks": sampler.summary(),
                "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                             "peak_source": "MEASURED_PEAKS.json hbm_gbs" if peaks else "fallback 6650",
                             "frac_of_8TBs_nominal": achieved / 8000.0, "bytes_per_element": alg / m.nels, "algorithmic_bytes_per_launch": alg,
                             "kernel": KERNEL_NAME.get(args.variant, "?") + (" (one launch per step)" if world == 1 else " (two launches per step: the CTAs owning interface rows on a high-priority stream with the exchange behind them, the interior CTAs on the ctx stream)"), "traffic": traffic,
                             "ncu_pipe_utilisation": pipes}}
        if check:
            line["check"] = check
        if e2e:
            h2d_all, d2h_all = (float(x) for x in io.tolist())
            line["e2e"] = {"value": nels_total / (e2e_ms * 1e-3), "unit": "elements/s", "h2d_bytes_per_step": int(h2d_all),
                           "d2h_bytes_per_step": int(d2h_all), "ms_per_step": e2e_ms,
                           "d2h_GBps_per_gpu": d2h_all / world / (e2e_ms * 1e-3) / 1e9, "numa_cpus_bound": numa_cpus}
        if world == 1 and not args.no_cpu:
            from oracle import pyoracle as orc
            orc.build()
            ne, sec = cpu_sample(args.cpu_n, 1)
            line["cpu_baseline"] = {"value": ne / sec, "unit": "elements/s", "cores": 1, "kind": "port",
                                    "sample": f"hex8 cube {args.cpu_n}^3 ({ne} elements) in {sec:.2f} s, oracle/ CPU "
                                              "restatement of src/main.F90:157-166; reference (Fortran+PETSc) not buildable here"}
    if da is not None:
        da.close()
    fx.close()
    if rank == 0:
        if world == 1 and not args.no_other:
            try:
                line["other_configs"] = other_configs(args, api, mg)
            except Exception as e:   # noqa: BLE001  -- the headline line must not depend on the side measurements
                line["other_configs"] = [{"error": repr(e)}]
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()
    return 0


def par_dev(ptr, n):
    from fem_b200.parallel import _DevArray
    return _DevArray(ptr, n)


if __name__ == "__main__":
    sys.exit(main())
