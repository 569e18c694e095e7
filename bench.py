#!/usr/bin/env python3
"""bench.py -- headline benchmark of the nsb200 hot path (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl nsb200|reference] [--n CELLS_PER_EDGE]

Workload (BASELINE.json configs[4], SURVEY.md 8 "C5"): 3D lid-driven cavity, Re = 100, IPCS
projection scheme with BDF-2, P2-P1 Taylor-Hood on the dolfin-style 73^3 x 6 tetrahedral unit
cube (9 934 793 DOF), dt = 1e-2, lid (1,0,0) on y = 1, no-slip elsewhere, start from rest.  One
"step" is one full IPCS time step: Newton on the tentative velocity (residual + Jacobian
assembly, Dirichlet rows, Jacobi-GMRES), pressure Poisson (rhs assembly, Jacobi-CG), velocity
correction (rhs assembly, Jacobi-CG on the mass matrix) and the BDF history shift.

Prints ONE JSON line (rank 0).  `value` = time steps per second with every input resident in HBM;
`e2e` = the same through the public solver API with the Dirichlet tables uploaded from pinned
host memory and the mixed solution vector copied back to the host inside the timed region.
`roofline` is for the kernel class with the largest share of the step, from CUDA-event pairs
recorded by libnsb200 around every launch on the launching stream (nsb_timer_*).
`cpu_baseline` / `--impl reference`: the CPU oracle (oracle/, plain C/OpenMP assembly and
Jacobi-GMRES/CG on all host threads, the "port" of the reference path -- FEniCS itself is
unavailable on this box) on a bounded sample of the same workload: the same cavity problem on a
16^3 mesh (112 724 DOF).  Its line reports the steps/s it MEASURED at that size under the size it
ran (`config.cells_per_edge`); the DOF-ratio extrapolation to the 73^3 workload is a separate,
clearly named field (`extrapolated_to_c5`), never the value.  The GPU line carries `same_size`:
both arms on the same 16^3 mesh, their ratio, and the relative error of the GPU fields against the
oracle's after the same steps (`parity_rel_err`); `checksums` (||u||_2, ||p - mean p||_2 after the K
timed steps, all-reduced over owned dofs) must agree across N = 1, 2, 4, 8.
"""
import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "time_steps_per_sec_ipcs_3d_cavity"
UNIT = "steps/s"
RE, DT = 100.0, 1.0e-2
CPU_SAMPLE_N = 16          # 16^3 x 6 tets, 112 724 DOF: a few seconds of CPU work per step


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="nsb200", choices=["nsb200", "reference"])
    ap.add_argument("--n", "--cells", dest="n", type=int, default=73, help="cells per edge of the unit cube (73 -> 9.93 M DOF)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--solver-option", action="append", default=[], metavar="KEY=VALUE",
                    help="IPCSSolver keyword for experiments, e.g. newton_forcing=0 (numbers only)")
    ap.add_argument("--ref-n", type=int, default=CPU_SAMPLE_N,
                    help="cells per edge of the bounded sample the CPU arm runs (and of the same-size GPU leg)")
    return ap.parse_args()


def mixed_dofs(n):
    """3 (2n+1)^3 + (n+1)^3"""
    return 3 * (2 * n + 1) ** 3 + (n + 1) ** 3


# ---------------------------------------------------------------------------------------------
# CPU arm: the oracle on a bounded sample
# ---------------------------------------------------------------------------------------------
def host_threads():
    """threads the CPU arm uses: every core this process may run on, set explicitly (torchrun exports
    OMP_NUM_THREADS=1, which would silently turn the all-core baseline into a single-thread one)"""
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return max(1, os.cpu_count() or 1)


def cpu_oracle_steps(n, steps, warmup, rtol=1e-12):
    """Time `steps` IPCS steps (after `warmup` untimed ones) of the CPU oracle on the n^3 cavity: C/OpenMP
    assembly + C/OpenMP Jacobi-GMRES(40)/CG at rtol 1e-12 (oracle/c/nsoracle.c) on all host threads.
    Returns (seconds per step, threads, oracle)."""
    import oracle
    threads = host_threads()
    om = oracle.mesh.hyper_rectangle((0.0, 0.0, 0.0), (1.0, 1.0, 1.0), (n, n, n))
    vel = [(w, None, (0.0, 0.0, 0.0)) for w in (1, 2, 3, 5, 6)] + [(4, None, (1.0, 0.0, 0.0))]
    orc = oracle.solvers.IPCSOracle(om, RE, vel, [], assembler="c", threads=threads)
    orc.linear_solver, orc.krylov_rtol = "krylov-c", rtol
    t, times = 0.0, []
    for s in range(warmup + steps):
        alpha = (1.0, -1.0, 0.0) if s == 0 else (1.5, -2.0, 0.5)
        t += DT
        t0 = time.perf_counter()
        orc.solve(alpha, DT, t)
        orc.advance_time()
        el = time.perf_counter() - t0
        if s >= warmup:
            times.append(el)
    return sum(times) / len(times), orc.asm.threads, orc


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    n = args.ref_n
    sec, threads, orc = cpu_oracle_steps(n, args.steps, args.warmup)
    value = 1.0 / sec                      # steps/s MEASURED on the n^3 sample -- nothing is scaled
    scale = mixed_dofs(n) / mixed_dofs(args.n)
    sample = ("%d timed IPCS steps (after %d warm-up steps) of the same cavity problem on the %d^3 x 6 mesh (%d DOF): "
              "C/OpenMP assembly + C/OpenMP Jacobi-GMRES(40)/CG rtol 1e-12 on %d threads, %.3f s/step; value is the "
              "rate measured at THIS size" % (args.steps, args.warmup, n, mixed_dofs(n), threads, sec))
    cfg = workload_config(n, 1)
    cfg["parallelism"] = "%d host threads (OpenMP), no GPU" % threads
    cfg["l2_policy"] = "n/a (CPU)"
    cfg["linear_solvers"] = "Jacobi-preconditioned GMRES(40) / CG, rtol 1e-12 (oracle/c/nsoracle.c nso_gmres, nso_pcg)"
    cfg["bounded_sample_of"] = workload_config(args.n, 1)["workload"]
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * sec,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": cfg,
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "extrapolated_to_c5": {"value": value * scale, "unit": UNIT, "dof_ratio": scale,
                               "assumption": "cost linear in DOF (favours the CPU: Krylov iteration counts grow with n); "
                                             "an estimate, not a measurement"},
        "newton_iterations": orc.newton_iterations[-args.steps:], "krylov_iterations": orc.krylov_iterations[-1],
        "checksums": oracle_checksums(orc),
        "note": "FEniCS/PETSc (the reference's own CPU path) is not installable on this box; kind=port is the "
                "CPU oracle restating it (oracle/c/nsoracle.c)",
    }
    print(json.dumps(line), flush=True)


def oracle_checksums(orc):
    import numpy as np
    u, p = orc.velocities[1], orc.old_pressure
    return {"u_l2": float(np.linalg.norm(u)), "p_minus_mean_l2": float(np.linalg.norm(p - p.mean()))}


def workload_config(n, gpus):
    return {"workload": "C5 3D lid-driven cavity IPCS/BDF-2, P2-P1 on %d^3x6 tets (%d DOF), Re=100, dt=1e-2"
                        % (n, mixed_dofs(n)),
            "cells_per_edge": n, "mixed_dofs": mixed_dofs(n),
            "parallelism": "1 GPU" if gpus == 1 else
                           "mesh partitioned into %d z-slabs, one rank per GPU (owned/ghost dofs, halo exchange before "
                           "every SpMV / assembly and all-reduced Krylov inner products, both over NVLink peer memory "
                           "[NCCL with NSB_HALO=nccl]); roofline and breakdown are rank 0's share" % gpus,
            "l2_policy": "inputs larger than L2 (Jacobian 6.5 GB, Krylov basis 3.3 GB per GPU at n=73)",
            "linear_solvers": "FP64 flexible GMRES(40), every Newton update to 1e-12 of the step's first residual except the "
                              "first update of a step (forcing term 0.001 x the Newton contraction observed at the previous "
                              "step, only when a second update is certain to follow: same Newton counts, same fields), "
                              "started from the previous step's update, right-preconditioned by a degree-10 Chebyshev polynomial "
                              "in D^-1 J that runs in single precision on the block-scalar part of J (mean block diagonals as "
                              "FP32, float4-padded vectors; operator, Arnoldi basis, residual test and solution stay FP64) "
                              "(velocity); persistent single-kernel CG rtol 1e-12 with a degree-8 Chebyshev polynomial "
                              "preconditioner (pressure) / CG rtol 1e-15 from u* + the previous step's correction, degree-2 single-precision Chebyshev polynomial preconditioner (mass)"}


# ---------------------------------------------------------------------------------------------
# clocks: sampled DURING the timed region through NVML
# ---------------------------------------------------------------------------------------------
class ClockSampler(threading.Thread):
    REASONS = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown", 0x4: "sw_power_cap",
               0x80: "hw_power_brake_slowdown"}

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.stop_flag = index, [], set(), False
        self.sm_max = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.sm_max = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def run(self):
        while not self.stop_flag and self.nv is not None:
            try:
                self.samples.append(self.nv.nvmlDeviceGetClockInfo(self.h, self.nv.NVML_CLOCK_SM))
                r = self.nv.nvmlDeviceGetCurrentClocksEventReasons(self.h) \
                    if hasattr(self.nv, "nvmlDeviceGetCurrentClocksEventReasons") \
                    else self.nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in self.REASONS.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            time.sleep(0.1)

    def result(self):
        self.stop_flag = True
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2] if s else None, "sm_max_mhz": self.sm_max, "reasons": sorted(self.reasons)}


# ---------------------------------------------------------------------------------------------
# GPU arm
# ---------------------------------------------------------------------------------------------
BIG_CLASSES = ("asm_jacobian", "asm_residual", "spmv_bsr", "spmv_bsr_f32")


def algorithmic_bytes(solver):
    """ALGORITHMIC bytes per launch of each kernel class (stated in DESIGN.md): FP64
    values, int32 indices, every value / vector entry moved once.  Sizes are those of the rank's
    share: the whole n^3 cavity on one GPU (nnz22 = 230 n^3 + 138 n^2 + 24 n + 1, App. C), the
    owned rows and local cells of the partition otherwise."""
    dev = solver.dev
    m = dev.mesh
    nverts, ncells = m.num_vertices, m.num_cells
    nnodes = dev.n_owned_nodes
    nnz22, nnz11 = dev.patterns["P2P2"].nnz, dev.patterns["P1P1"].nnz
    nV, nQ = 3 * nnodes, dev.n_owned_verts
    geo = 8 * 3 * nverts + 4 * 4 * ncells + 4 * 10 * ncells    # coords + cell->vertex + cell->node maps
    return {
        "spmv_bsr": 8 * 9 * nnz22 + 4 * nnz22 + 16 * nV + 4 * nnodes,
        # single-precision SpMV of the preconditioner on the block-scalar part of J: one float + one column
        # index per block entry, x and y as one float4 per node, row offsets
        # + the Chebyshev step fused into its epilogue: r and z read + written, d_new written (float4 each), D^-1;
        # the gathered copy of x is its bf16 shadow (8 B per node read, 8 B written for d_new), x itself is read once
        # (own row, float4)
        "spmv_bsr_f32": 4 * nnz22 + 4 * nnz22 + 4 * nnodes + (16 + 8 + 8 + 32 + 32 + 16 + 24) * nnodes,
        "spmv_csr": None,                                       # two different matrices share the class
        "asm_jacobian": 8 * 9 * nnz22 + geo + 8 * nV,
        "asm_residual": 8 * nV + geo + 8 * (3 * nV + nQ),
    }, dict(nV=nV, nQ=nQ, nnz22=nnz22, nnz11=nnz11)


def make_cavity_solver(n, rank=0, world=1, local=0, **solver_options):
    """the C5 problem on the n^3 x 6 unit cube: lid (1,0,0) on y = 1, no-slip elsewhere, from rest"""
    import nsb200
    V = nsb200.VelocityBCType
    mesh, markers = nsb200.hyper_cube(3, n)
    part = None
    if world > 1:
        # mesh-partitioned run: z-slabs, one rank per GPU, halos + all-reduces in libnsb200
        from nsb200 import partition as npart
        part = npart.connect_distributed(npart.build_partition(mesh, rank, world))
    ts = nsb200.BDFTimeStepping(0.0, 1.0e6, order=2, desired_start_time_step=DT)
    solver = nsb200.IPCSSolver(mesh, markers, "standard", ts, device_index=local, partition=part, **solver_options)
    solver.set_equation_coefficients(dict(convective_term=1.0, pressure_term=1.0, viscous_term=1.0 / RE))
    solver.set_boundary_conditions([(V.no_slip, w, None) for w in (1, 2, 3, 5, 6)] + [(V.constant, 4, (1.0, 0.0, 0.0))])
    return solver, ts


def same_size_leg(n_ref, steps, local, **solver_options):
    """Both arms on the SAME n_ref^3 mesh for the same `steps` time steps from rest: measured steps/s of each
    (the GPU's through the public API incl. the per-step solution copy to the host), their ratio, and the
    relative max-norm error of the GPU fields against the oracle's (pressure up to its mean: the all-Neumann
    problem fixes p only up to a constant)."""
    import numpy as np
    import torch
    solver, ts = make_cavity_solver(n_ref, local=local, **solver_options)

    def run(k):
        for _ in range(k):
            ts.update_coefficients()
            solver.solve(upload_bcs=True)
            solver.fetch_solution()
            ts.advance_time()
            solver.advance_time()
    state = solver.get_state()
    run(steps)                                  # untimed pass: one-time setup (constant operators, pinned buffers)
    solver.set_state(state)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    run(steps)
    torch.cuda.synchronize()
    gpu_sec = (time.perf_counter() - t0) / steps
    u = solver._velocities[1].cpu().numpy()
    p = solver._old_pressure.cpu().numpy()
    cpu_sec, threads, orc = cpu_oracle_steps(n_ref, steps, 0)          # timed: tolerances matched to the GPU arm (1e-12)
    # untimed: the parity checker, three digits tighter (measured on the oracle itself, 16^3, against its 3e-16 run:
    # at 1e-14 its own fields are 9e-10 / 8e-10 off, at 1e-15 1e-10 / 6e-11)
    _, _, chk = cpu_oracle_steps(n_ref, steps, 0, rtol=1e-15)
    ou, op = chk.velocities[1], chk.old_pressure
    err = {"velocity": float(np.abs(u - ou).max() / np.abs(ou).max()),
           "pressure_minus_mean": float(np.abs((p - p.mean()) - (op - op.mean())).max() / np.abs(op - op.mean()).max())}
    out = {"cells_per_edge": n_ref, "mixed_dofs": mixed_dofs(n_ref), "steps": steps,
           "gpu_steps_per_s_e2e": 1.0 / gpu_sec, "cpu_steps_per_s": 1.0 / cpu_sec, "cpu_threads": threads,
           "ratio": cpu_sec / gpu_sec, "parity_rel_err": err,
           "newton_iterations": {"gpu": solver.newton_iterations[-steps:], "cpu": orc.newton_iterations[-steps:]},
           "parity_checker": "the same oracle at rtol 1e-15 (untimed; its own error at that tolerance: 1e-10)",
           "checksums": {"gpu": solver.checksums(lagged=True), "cpu": oracle_checksums(chk)}}
    solver.dev.close()
    return out, cpu_sec, threads


def run_nsb200(args):
    import torch
    import __graft_entry__
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py --impl nsb200 needs a CUDA device (no CPU fallback)")
    torch.cuda.set_device(local)
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("cpu:gloo,cuda:nccl", device_id=torch.device("cuda", local))
    if local == 0:
        __graft_entry__.build()
    if world > 1:
        dist.barrier()
    import nsb200
    n = args.n
    opts = {k: float(v) if "." in v or "e" in v else int(v) for k, v in (o.split("=", 1) for o in args.solver_option)}
    solver, ts = make_cavity_solver(n, rank, world, local, **opts)
    dev = solver.dev

    def step(e2e=False):
        ts.update_coefficients()
        solver.solve(upload_bcs=e2e)
        h2d = solver.h2d_bytes
        d2h = 0
        if e2e:
            # the device -> host copy of this step's solution overlaps the next step (side stream); the timed region
            # ends with a device-wide synchronize, so every copy it started has arrived inside it
            _, d2h = solver.fetch_solution(wait=False)
        ts.advance_time()
        solver.advance_time()
        return h2d, d2h

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(K, e2e):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        h2d = d2h = 0
        for _ in range(K):
            a, b = step(e2e)
            h2d, d2h = a, b
        if e2e:
            solver.wait_solution()              # the last step's copy has arrived before the clock stops
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item()), h2d, d2h

    for w in range(args.warmup):
        step(e2e=(w == args.warmup - 1))    # the last warm-up step goes through the end-to-end path once, so that its
        #                                     one-time pinned allocations and map uploads are not inside a timed region
    sampler = ClockSampler(local)
    sampler.start()
    l0 = solver.launch_count()
    state = solver.get_state()              # all three legs advance the SAME K time steps
    # leg 1, the headline: CUDA-event pairs only around the large kernels (assembly, velocity SpMVs: ~170 launch
    # groups per step) -- the roofline numbers come from here
    solver.timer_enable(True, classes=BIG_CLASSES)
    ms, _, _ = timed(args.steps, False)
    big = solver.timer_report()
    solver.timer_enable(False)
    launches = solver.launch_count() - l0
    clocks = sampler.result()
    # leg 2, the breakdown: every launch group of every class timed (an event pair costs about a microsecond on
    # each of the ~1400 small Krylov launches per step, so this leg is not the headline)
    solver.set_state(state)
    solver.timer_enable(True)
    ms_instrumented, _, _ = timed(args.steps, False)
    classes = solver.timer_report()
    solver.timer_enable(False)
    # leg 3, end to end through the public API: host Dirichlet tables in, mixed solution out, every step
    solver.set_state(state)
    del state
    ms_e2e, h2d, d2h = timed(args.steps, True)
    checksums = solver.checksums(lagged=True)      # collective on a partitioned run: before the tear-down
    if world > 1:
        # collective tear-down BEFORE rank 0 goes off to format the result: every rank leaves together
        per_rank = [None] * world
        dist.all_gather_object(per_rank, dict(cells=dev.mesh.num_cells, owned_nodes=dev.n_owned_nodes,
                                              ghost_nodes=dev.spaces.n_nodes - dev.n_owned_nodes))
        barrier()
        dev.close()                      # collective: ncclCommDestroy of libnsb200's communicator
        dist.destroy_process_group()
    else:
        per_rank = None
    if rank != 0:
        return
    K = args.steps
    value = K / (ms * 1e-3)                 # N > 1: ONE partitioned problem advanced by all ranks together
    e2e_value = K / (ms_e2e * 1e-3)
    alg, sizes = algorithmic_bytes(solver)
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            peaks = json.load(f)
    except (OSError, ValueError):
        peaks = {}
    hbm_peak, peak_src = (peaks["hbm_gbs"], "measured (MEASURED_PEAKS.json hbm_gbs)") if "hbm_gbs" in peaks \
        else (6650.0, "fallback (B200_PROFILING.md)")
    total_ms = sum(v[0] for v in classes.values())
    breakdown = {k: {"ms_per_step": v[0] / K, "launch_groups_per_step": v[1] / K,
                     "share_of_kernel_time": v[0] / total_ms if total_ms else None} for k, v in classes.items()}
    dom = max((k for k in BIG_CLASSES if alg.get(k)), key=lambda k: big[k][0])
    dur_ms = big[dom][0] / max(big[dom][1], 1)
    achieved = alg[dom] / (dur_ms * 1e-3) / 1e9
    # dram__bytes_read.sum + dram__bytes_write.sum of one launch of the same kernel on the same workload,
    # from the committed `ncu --set full` capture (profiles/): only valid for the full C5 mesh on one GPU
    traffic = None
    if n == 73 and world == 1:
        try:
            with open(os.path.join(ROOT, "profiles", "r2_ncu_full_summary.json")) as f:
                traffic = json.load(f)[dom]["traffic_bytes"]
        except (OSError, ValueError, KeyError):
            traffic = None
    roofline = {"kernel": dom, "bound": "hbm", "achieved": achieved, "peak": hbm_peak, "unit": "GB/s",
                "frac": achieved / hbm_peak, "traffic": traffic, "peak_source": peak_src,
                "algorithmic_bytes_per_launch": alg[dom], "avg_launch_ms": dur_ms,
                "launches_timed": big[dom][1]}
    for k in BIG_CLASSES:
        if big[k][1] and alg.get(k):
            d_ms = big[k][0] / big[k][1]
            breakdown[k]["GB/s"] = alg[k] / (d_ms * 1e-3) / 1e9
            breakdown[k]["frac_of_hbm_peak"] = breakdown[k]["GB/s"] / hbm_peak
    # element kernels: the FP64 view next to the HBM view (SURVEY.md 8d "report both fractions"): share of cycles the
    # FP64 pipe was busy, from the committed ncu capture of the same kernel on the same workload, and the FP64 FMA peak
    # measured live by the library (nsb_fp64_peak) -- neither assembly kernel is bound by HBM or by the FP64 pipe
    if n == 73 and world == 1:
        try:
            with open(os.path.join(ROOT, "profiles", "r2_ncu_full_summary.json")) as f:
                summ = json.load(f)
            for k in ("asm_jacobian", "asm_residual"):
                m = summ[k]["metrics"]
                breakdown[k]["fp64_pipe_active_pct_ncu"] = float(m["sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active"].split()[0])
                breakdown[k]["l1tex_throughput_pct_ncu"] = float(m["l1tex__throughput.avg.pct_of_peak_sustained_elapsed"].split()[0])
        except (OSError, ValueError, KeyError):
            pass
    try:
        fp64_peak = dev.fp64_peak()
    except RuntimeError:
        fp64_peak = None
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": args.warmup,
        "ms_per_step": ms / K, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic", "config": workload_config(n, world), "clocks": clocks,
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h},
        "gpu_launches": launches, "roofline": roofline, "breakdown": breakdown,
        "breakdown_note": "ms_per_step / shares from a second pass over the same %d steps with every kernel class "
                          "event-timed (%.1f ms/step); GB/s of the large kernels and the roofline from the headline pass"
                          % (K, ms_instrumented / K),
        "newton_iterations": solver.newton_iterations[-K:], "newton_iterations_all_steps": solver.newton_iterations,
        "krylov_iterations": solver.krylov_iterations[-1],
        "chebyshev_fallbacks": solver.chebyshev_fallbacks,
        "checksums": checksums, "fp64_fma_peak_tflops_measured": fp64_peak,
        "assembly_gdof_per_s": {k: sizes["nV"] / (big[k][0] / big[k][1] * 1e-3) / 1e9
                                for k in ("asm_jacobian", "asm_residual") if big[k][1]},
    }
    if world == 1 and not args.no_cpu_baseline:
        ss, sec, threads = same_size_leg(args.ref_n, 2, local, **opts)
        scale = mixed_dofs(args.ref_n) / mixed_dofs(n)
        line["cpu_baseline"] = {
            "value": 1.0 / sec, "unit": UNIT, "cores": threads, "kind": "port",
            "sample": "2 IPCS steps from rest of the same cavity problem on the %d^3 x 6 mesh (%d DOF): C/OpenMP assembly + "
                      "C/OpenMP Jacobi-GMRES(40)/CG rtol 1e-12 on %d threads, %.3f s/step; value is the rate MEASURED at "
                      "that size (DOF-ratio estimate for the %d^3 workload: %.3g steps/s, not a measurement)"
                      % (args.ref_n, mixed_dofs(args.ref_n), threads, sec, n, scale / sec)}
        line["same_size"] = ss
    if per_rank is not None:
        line["partition"] = per_rank
        line["halo_transport"] = solver.dev.halo_transport
    print(json.dumps(line), flush=True)


def main():
    args = parse()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_nsb200(args)


if __name__ == "__main__":
    main()
