#!/usr/bin/env python
"""bench.py -- elements/s of the (Ke + f_int) global assembly hot path on B200.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--n CUBE] [--impl reference]

Workload (BASELINE.json configs[1]): synthetic structured hex8 cube n^3 (default 100^3 =
1M elements, 3.09M dof, 245M nonzeros) per GPU, linear elastic, jittered interior nodes so
elements are not congruent.  A step = one femx_assemble_K_fint call: every CSR value of K and
the internal force f = sum_e Ke u_e are (re)computed from coordinates and u.  `value` is
device-timed with inputs resident in HBM; `e2e` goes through the same C-ABI call with pinned
HOST buffers (u H2D, K values + f D2H inside the timed region).
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
KERNEL_NAME = {0: "hex_sym6_kernel", 1: "elem_rows_kernel + gather_rows_kernel", 2: "hex_fused_kernel", 3: "hex_ws_kernel",
               4: "hex_sym5_kernel", 5: "hex_sym2_kernel", 6: "hex_sym3_kernel", 7: "hex_sym4_kernel", 8: "hex_sym5_kernel", 12: "hex_sym6_kernel", 13: "hex_sym7_kernel"}   # conn 32 + mat 4 + coords 24.73 + u 24.73 + K 1963.5 + R 24.73


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
    from fem_b200 import mesh as fm, meshgen as mg
    if nranks == 1:
        m = mg.jitter(mg.hex_cube(n, clamp=False))
    else:
        # weak scaling: a column of nranks cubes, one n^3 block per rank (z-slabs); each rank
        # assembles the rows of its own elements (SURVEY.md 8e partitioning by element)
        m = mg.hex_block(n, n, n, z0=rank * n, nz_total=n * nranks)
    return m, mg.displacement_field(m)


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
    ap.add_argument("--tile", type=int, default=0)
    ap.add_argument("--lz", type=int, default=0)
    ap.add_argument("--dbg", type=int, default=0, help="diagnostic: skip parts of the fused kernel (results invalid)")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)

    import torch
    import torch.distributed as dist
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (femx has no CPU fallback)")
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    from fem_b200 import api
    m, u = build_workload(args.n, world, rank)
    fx = api.Femx(device=local_rank, quirks=api.Q_REFERENCE, assembly_variant=args.variant, tile=args.tile, lz=args.lz, dbg=args.dbg,
                  peer_in_kernel=args.exchange == "peer-fused")
    t0 = time.time()
    fx.set_mesh(m).build_plan()
    plan_s = time.time() - t0
    nnz, nrows = fx.nnz, fx.nrows

    d_u = fx.device_buffer(u.nbytes).upload(u)
    d_f = fx.device_buffer(8 * nrows)
    da = None
    if world > 1:
        # interface rows go to their PETSc owners over NCCL every step (fem_b200/parallel.py)
        import ctypes as C
        from fem_b200 import parallel as par
        da = par.DistributedAssembly(fx, dist, torch.device("cuda", local_rank), with_merge=False, peer=args.exchange != "nccl")
        vals_ptr = C.c_void_p(da.vals_local.data_ptr())
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    else:
        d_vals = fx.device_buffer(8 * nnz)
        vals_ptr = d_vals.ptr

    def step_device():
        fx._ck(fx.L.femx_assemble_K_fint(fx.h, d_u.ptr, vals_ptr, d_f.ptr, api.DEVICE))
        ms = fx.last_kernel_ms
        if da is not None and not da.peer:   # peer mode: the kernel stored the interface rows into the owners' arenas
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

    sampler = ClockSampler(local_rank)
    sampler.start()                      # samples from warm-up to the end of the e2e loop: the GPU is under load throughout
    for _ in range(args.warmup):
        step_device()
    barrier()
    l0 = fx.launches
    w0 = time.perf_counter()
    dev_ms = 0.0
    for _ in range(args.steps):
        dev_ms += step_device()
    barrier()
    wall_ms = 1e3 * (time.perf_counter() - w0)
    launches = fx.launches - l0

    # e2e: pinned host buffers through the same C-ABI call
    e2e = None
    if not args.no_e2e:
        hu = fx.pinned(u.size); hu.array[:] = u.reshape(-1)
        hv = fx.pinned(nnz); hf = fx.pinned(nrows)

        def step_host():
            fx._ck(fx.L.femx_assemble_K_fint(fx.h, hu.ptr, hv.ptr, hf.ptr, api.HOST))

        step_host()
        barrier()
        e0 = time.perf_counter()
        ke = max(2, min(args.steps, 5))
        for _ in range(ke):
            step_host()
        barrier()
        e2e_ms = 1e3 * (time.perf_counter() - e0) / ke
        # parity spot check of what came back (size-independent properties; tests/ hold the oracle parity)
        assert np.isfinite(hv.array[:1000]).all() and np.isfinite(hf.array).all()
        e2e = {"ms": e2e_ms, "h2d": int(u.nbytes), "d2h": int(8 * nnz + 8 * nrows)}
        hu.free(); hv.free(); hf.free()
    sampler.stop_flag = True
    sampler.join(timeout=2)

    t = torch.tensor([dev_ms, wall_ms, e2e["ms"] if e2e else 0.0], device="cuda", dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dev_ms, wall_ms, e2e_ms = (float(x) for x in t.tolist())
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
            tj = json.load(open(os.path.join(ROOT, "profiles", "r01_traffic.json")))
            if args.n == 100:
                ent = tj.get(KERNEL_NAME.get(args.variant, "?"), {})
                traffic = ent.get("traffic_bytes_per_launch")
                if "lsu_data_pipe_pct_of_peak" in ent:   # what actually bounds the kernel (committed ncu capture, not measured live)
                    pipes = {k: ent[k] for k in ("lsu_data_pipe_pct_of_peak", "fp64_pipe_pct_of_peak", "dram_pct_of_peak")}
        except Exception:
            pass
        line = {"metric": "elements/sec (Ke+f_int assembly)", "value": value, "unit": "elements/s", "n_gpus": world,
                "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                "config": {"workload": f"hex8 cube {args.n}^3 per GPU, jittered, K + f_int (BASELINE configs[1])",
                           "elements_per_gpu": m.nels, "dof_per_gpu": int(m.pdim * m.nlnds), "nnz_per_gpu": int(nnz),
                           "l2": "outputs (8*nnz bytes) exceed the 126 MB L2; no flush needed",
                           "parallelism": (f"z-slab element partition x{world}, PETSc row split, interface rows "
                                           + ((("stored by the assembly kernel" if args.exchange == "peer-fused" else "pushed by femx's peer kernel")
                                               + " into the owner's arena over NVLink + device-side step barrier")
                                              if da.peer else "sent with NCCL all_to_all after the kernel")
                                           + f" ({int(da.ep.recv_counts.sum()) * 8} B received on rank 0 per step)") if world > 1 else "1 GPU",
                           "variant": args.variant, "tile": args.tile, "lz": args.lz, "plan_build_s": round(plan_s, 3), "wall_ms_per_step": wall_ms / args.steps},
                "gpu_launches": int(launches),
                "clocks": sampler.summary(),
                "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                             "peak_source": "MEASURED_PEAKS.json hbm_gbs" if peaks else "fallback 6650",
                             "frac_of_8TBs_nominal": achieved / 8000.0, "bytes_per_element": alg / m.nels, "algorithmic_bytes_per_launch": alg,
                             "kernel": KERNEL_NAME.get(args.variant, "?") + " (one launch per step)", "traffic": traffic,
                             "ncu_pipe_utilisation": pipes}}
        if e2e:
            line["e2e"] = {"value": nels_total / (e2e_ms * 1e-3), "unit": "elements/s", "h2d_bytes_per_step": e2e["h2d"],
                           "d2h_bytes_per_step": e2e["d2h"], "ms_per_step": e2e_ms}
        if world == 1 and not args.no_cpu:
            from oracle import pyoracle as orc
            orc.build()
            ne, sec = cpu_sample(args.cpu_n, 1)
            line["cpu_baseline"] = {"value": ne / sec, "unit": "elements/s", "cores": 1, "kind": "port",
                                    "sample": f"hex8 cube {args.cpu_n}^3 ({ne} elements) in {sec:.2f} s, oracle/ CPU "
                                              "restatement of src/main.F90:157-166; reference (Fortran+PETSc) not buildable here"}
        print(json.dumps(line))
    if da is not None:
        da.close()
    fx.close()
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
