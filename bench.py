#!/usr/bin/env python
"""bench.py -- BASELINE.json's metric on BASELINE.json's config.

Workload (N=1): configs[1], "3-D Taylor-Green vortex, 16^3 elements, ngl=4, one
B200".  A *step* is one RK-5BS time step of the vorticity-transport equation
through the reference's own time loop (run_case.py:152-162 -> TS -> evalRHS,
base_problem.py:111-136): 7 right-hand-side evaluations (FSAL), each = BC scatter
+ KLE solve (CG on K to --rtol; preconditioner --pc: the two-level element Schwarz
method by default, node-block Jacobi in the "alt_pc_bjacobi" block) + SpMV with Rw,
Krhs, SrT, DivSrT, Curl + the pointwise stress kernel.  The operators are assembled once
before the loop, as in the reference (setUpSolver, base_problem.py:156-179); the
assembly is timed separately and reported under "assembly".

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]

value  = velocity DOFs advanced one time step per second, fields resident in HBM;
e2e    = the same through the public API with the fields living on the HOST:
         every step uploads vorticity+velocity from pinned host memory and reads
         them back (what the reference's post-step viewer does), BC values are
         evaluated on the host and scattered every RHS evaluation.
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "kle_timestep_throughput"
UNIT = "DOF/s"          # velocity dofs advanced one RK5BS time step per second


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--nelem", type=int, default=16)
    ap.add_argument("--ngl", type=int, default=4)
    ap.add_argument("--rtol", type=float, default=1e-10, help="KSP relative tolerance")
    ap.add_argument("--dt", type=float, default=1e-3, help="fixed time step (-ts_adapt_type none)")
    ap.add_argument("--pc", default="asm", choices=["jacobi", "bjacobi", "none", "asm"],
                    help="asm: element-block additive Schwarz with a Q1 coarse level (--coarse Chebyshev steps)")
    ap.add_argument("--coarse", type=int, default=5, help="-pc_asm_coarse (0: one-level Schwarz)")
    ap.add_argument("--no-alt", action="store_true", help="skip the comparison pass with -pc_type bjacobi (round-1 default)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-matfree", action="store_true", help="skip the extra matrix-free (-kle_matfree 1) pass")
    ap.add_argument("--no-asm-warm", action="store_true", help="skip the warm re-assembly passes (uniform + perturbed mesh)")
    ap.add_argument("--fischer", type=int, default=16,
                    help="-ksp_fischer_guess 1,<n>: initial guess from the last n solutions (0: off, PETSc default)")
    ap.add_argument("--opt", action="append", default=[], metavar="NAME=VALUE",
                    help="extra options-database entry, e.g. --opt ksp_cg_single_reduction=1 (repeatable)")
    ap.add_argument("--tune", action="append", default=[], metavar="KEY=INT",
                    help="pyn_tune(key, value) before the run, e.g. --tune halo_wait_at_start=1 (repeatable)")
    ap.add_argument("--check-every", type=int, default=0, help="KSP status poll interval (0: library default)")
    return ap.parse_args()


def load_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs, copy burst)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(index), f"--query-gpu={self.Q}",
                                       "--format=csv,noheader,nounits", "-lms", "100"],
                                      stdout=self.f, stderr=subprocess.DEVNULL)
        except OSError:
            self.p = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": []}
        if self.p is None:
            return out
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except Exception:
            self.p.kill()
        self.f.flush()
        self.f.seek(0)
        sm, smax, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in self.f.read().splitlines():
            c = [x.strip() for x in line.split(",")]
            if len(c) < 6:
                continue
            try:
                sm.append(float(c[0]))
                smax.append(float(c[1]))
            except ValueError:
                continue
            for n, v in zip(names, c[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        self.f.close()
        os.unlink(self.f.name)
        if sm:
            out = {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(smax)), "reasons": sorted(reasons),
                   "samples": len(sm)}
        return out


def workload_name(args):
    which = "BASELINE.json configs[1]" if (args.nelem, args.ngl) == (16, 4) else "NOT the configs[1] size 16^3 / ngl=4"
    return (f"3-D Taylor-Green vortex, {args.nelem}^3 elements, ngl={args.ngl}, unit box "
            f"({which}); one RK5BS time step = 7 evalRHS")


def case_config(nelem, ngl):
    """cases/taylor-green3d.yaml (the reference's src/cases/taylor-green3d.yaml) with the box set
    to BASELINE.json configs[1]"""
    from pynama_b200.run_case import loadCase
    cfg = loadCase("taylor-green3d")
    cfg["domain"]["ngl"] = ngl
    cfg["domain"]["box-mesh"]["nelem"] = [nelem] * 3
    cfg["time-solver"]["max-steps"] = 10 ** 9
    cfg["time-solver"]["end-time"] = 1e9
    return cfg


# ---------------------------------------------------------------------------------------------
# B200 arm
# ---------------------------------------------------------------------------------------------
def build_problem(args, pc=None):
    from pynama_b200.cases.base_problem import BaseProblem
    from pynama_b200.petsc import Options
    pc = pc or args.pc
    o = Options()
    o.setValue("ksp_type", "cg")
    o.setValue("pc_type", pc)
    o.setValue("pc_asm_coarse", args.coarse if pc == "asm" else 0)
    o.setValue("ksp_rtol", args.rtol)
    o.setValue("ksp_max_it", 50000)
    if args.fischer > 0:
        o.setValue("ksp_fischer_guess", f"1,{args.fischer}")
    o.setValue("ts_adapt_type", "none")
    o.setValue("ts_dt", args.dt)
    if (int(os.environ.get("WORLD_SIZE", "1")) > 1 and pc != "asm"
            and not any(kv.startswith("ksp_cg_single_reduction") for kv in args.opt)):
        # row-partitioned runs: CG with one synchronisation point per iteration (PETSc's
        # -ksp_cg_single_reduction); with the peer-memory back end the halo exchange is fused too
        args.opt.append("ksp_cg_single_reduction=1")
    for kv in args.opt:
        name, _, val = kv.partition("=")
        o.setValue(name, val or "1")
    fem = BaseProblem(case_config(args.nelem, args.ngl), case="taylor-green3d")
    fem.setUp()
    t0 = time.perf_counter()
    fem.setUpSolver()
    fem.setup_wall_s = time.perf_counter() - t0
    fem.ts._post = None            # the post-step viewer is file output (out of scope); e2e reads back instead
    fem.ts.setSolution(fem.vort)
    return fem


def bc_bytes_per_rhs(fem):
    """host->device bytes of one evalRHS: merged BC index + value arrays of both fields"""
    bc = fem.dom.getBoundaryConditions()
    n = 0
    for name in ("vorticity", "velocity"):
        idx, vals = bc.getValuesMerged(name, 0.0, fem.nu)
        n += np.asarray(idx).size * 8 + np.asarray(vals).size * 8
    return n


def run_b200(args, rank, world):
    from pynama_b200 import _lib
    from pynama_b200.petsc import Context
    ctx = Context.get()
    if args.check_every > 0:
        _lib.call("pyn_tune", b"ksp_check_every", args.check_every)
    for kv in args.tune:
        key, _, val = kv.partition("=")
        _lib.call("pyn_tune", key.encode(), int(val or "1"))
    FP64_PEAK["TFLOPs"] = ctx.fp64_peak()          # DFMA chains, measured on this device before anything else runs
    ctx.prof_enable(True)
    fem = build_problem(args)
    asm_kernels = {k: ctx.prof_get(k) for k in ("elem_kle", "numeric")}
    ctx.prof_enable(False)
    dm = fem.dom.getDeviceMesh()
    K = fem.mat.K
    ndof = K.getSize()[0]
    nvort = fem.vort.getSize()
    asm = {p: dm.phase_ms(p) for p in ("symbolic", "elem_kle", "kle_numeric", "elem_ops", "ops_numeric")}

    def one_step():
        fem.ts.step()

    # Every timed region below does the SAME work: from the initial condition (state kept on the
    # device), with an empty Fischer basis, W untimed warm-up steps and then the K timed steps.
    # (The Krylov iteration count drifts as the vortex evolves, and a basis that has already seen
    # the timed steps would make a repeated region cheaper than the first.)
    vel = fem.solverKLE.getSolution()
    snap = {"vort": fem.vort.copy(), "vel": vel.copy(), "time": fem.ts.time, "step": fem.ts.step_number}

    def prepare(warmup=None):
        snap["vort"].copy(fem.vort)
        snap["vel"].copy(vel)
        fem.ts.time, fem.ts.step_number, fem.ts._fsal_valid = snap["time"], snap["step"], False
        if args.fischer > 0:
            fem.solverKLE.solver.setUseFischerGuess(1, args.fischer)      # drops the basis
        for _ in range(args.warmup if warmup is None else warmup):
            one_step()
        ctx.sync()

    # checksum: norms of the fields CHECK_STEPS steps after the initial condition -- the same number at
    # every N (SCALE shows the N-rank runs computed the same fields) and in the CPU arm (cpu_baseline /
    # --impl reference run the same steps on the same mesh)
    prepare(CHECK_STEPS)
    checksum = dict(field_checksum(fem.vort.getGlobalArray(), vel.getGlobalArray()), after_steps=CHECK_STEPS)

    restore = prepare
    prepare()
    barrier(world)
    # timed region A: K steps, nothing but the work -> value
    sampler = ClockSampler(ctx.device)
    l0 = ctx.launch_count()
    ctx.sync()
    ctx.timer_start()
    for _ in range(args.steps):
        one_step()
    ms = ctx.timer_stop()
    ctx.sync()
    barrier(world)
    clocks = sampler.stop()
    launches = ctx.launch_count() - l0
    # timed region B: the same K steps again with a CUDA-event pair around every launch of the
    # hot kernel classes (costs ~25 us per Krylov iteration, hence not in region A) -> roofline
    restore()
    its0, solves0, evals0 = fem.kspIterations, fem.kspSolves, fem.ts.rhs_evals
    ctx.prof_enable(True)
    ctx.sync()
    ctx.timer_start()
    for _ in range(args.steps):
        one_step()
    ms_prof = ctx.timer_stop()
    ctx.sync()
    barrier(world)
    n_spmv, ms_spmv = ctx.prof_get("spmv_ksp")
    n_oth, ms_oth = ctx.prof_get("spmv")
    n_upd, ms_upd = ctx.prof_get("cg_update")
    n_dir, ms_dir = ctx.prof_get("cg_direction")
    ctx.prof_enable(False)
    its = fem.kspIterations - its0
    solves = fem.kspSolves - solves0
    evals = fem.ts.rhs_evals - evals0
    ms = max_over_ranks(ms, world)
    ms_per_step = ms / args.steps

    # dominant kernel: SpMV(K) fused with p.Ap inside CG
    stored, spmv_bytes = K.bytes()
    nnz_ref, nblocks = K.nnz()
    peak, peak_src = load_peaks()
    # CG launches iterations in chunks of 8 and polls a pinned status word; launches issued after
    # convergence return at once.  Only the `its` live iterations move bytes, so the average is
    # taken over those (the few-us dead launches stay in the numerator: slightly pessimistic).
    live = max(min(its, n_spmv), 1)
    avg_ms = ms_spmv / live
    achieved = spmv_bytes / (avg_ms * 1e-3) / 1e9 if n_spmv else 0.0
    per_rank = None
    if world > 1:
        # every rank multiplies its own rows: bytes = the blocks of the OWNED rows (pyn_mat_bytes), time =
        # that rank's launches; the slowest rank is the one reported, the spread is printed beside it
        lo = -float(ctx.allreduce([-achieved], "max")[0])
        hi = float(ctx.allreduce([achieved], "max")[0])
        per_rank = {"rank0": round(achieved, 1), "min": round(lo, 1), "max": round(hi, 1),
                    "bytes_rank0": int(spmv_bytes),
                    "bytes_max": int(ctx.allreduce([float(spmv_bytes)], "max")[0]),
                    "avg_launch_us_max": round(float(ctx.allreduce([avg_ms * 1e3], "max")[0]), 2)}
        achieved = lo
    roofline = {"bound": "hbm", "kernel": "spmv_kernel<3,3,32,DOT> on K (row-plane BSR)",
                "achieved": round(achieved, 1), "peak": peak, "unit": "GB/s", "frac": round(achieved / peak, 4),
                "peak_source": peak_src, "bytes_per_launch": int(spmv_bytes), "launches": int(live),
                "dead_launches": int(n_spmv - live),
                "avg_launch_us": round(avg_ms * 1e3, 2), "share_of_step": round(ms_spmv / ms_prof, 4),
                "measured_in": "second pass of the same K steps with per-launch CUDA events "
                               f"({ms_prof / args.steps:.1f} ms/step with the events on)",
                "frac_of_nominal_8TBs": round(achieved / 8000.0, 4),
                "reference_equivalent_GBs": round((12 * nnz_ref + 16 * ndof + 8 * ndof) / (avg_ms * 1e-3) / 1e9, 1)
                if n_spmv else 0.0,
                "traffic": read_traffic() if world == 1 and args.nelem == 16 and args.ngl == 4 else None}
    if per_rank:
        roofline["over_ranks"] = per_rank

    # the next kernels of the step, each against the roofline that bounds it
    nloc = (K.getLocalSize()[0] if hasattr(K, "getLocalSize") else ndof)
    upd_bytes = 8 * nloc * 8 + (nloc // 3) * 9 * 8        # x, r, z, p, w read / written (8 passes) + the 3x3 PC blocks
    upd_us = ms_upd / max(n_upd, 1) * 1e3
    other_rooflines = [
        {"kernel": "cg_update_kernel (x, r, z update + node-block Jacobi + r.z, z.z + convergence test)",
         "bound": "L2 / launch latency (vectors are L2-resident: 2.8 MB each)", "bytes_per_launch": int(upd_bytes),
         "avg_launch_us": round(upd_us, 2), "achieved_GBs": round(upd_bytes / (upd_us * 1e-6) / 1e9, 1) if n_upd else None,
         "frac_of_hbm_peak": round(upd_bytes / (upd_us * 1e-6) / 1e9 / peak, 4) if n_upd else None},
    ]

    e2e = None
    if not args.no_e2e:
        e2e = run_e2e(args, fem, ctx, world, ndof, nvort, restore)

    # the same steps with the products K*x of the Krylov loop evaluated matrix-free
    # (-kle_matfree 1): reported beside the stored-matrix path, not instead of it
    matfree = None
    if not args.no_matfree and args.ngl <= 6:
        K.setMatFree(True)
        restore()
        barrier(world)
        its0 = fem.kspIterations
        ctx.timer_start()
        for _ in range(args.steps):
            one_step()
        ms_mf = max_over_ranks(ctx.timer_stop(), world) / args.steps
        its_mf = (fem.kspIterations - its0) / args.steps
        ctx.prof_enable(True)
        one_step()
        ctx.sync()
        n_mf, t_mf = ctx.prof_get("spmv_ksp")
        ctx.prof_enable(False)
        matfree = {"option": "-kle_matfree 1", "value": ndof / (ms_mf * 1e-3), "unit": UNIT, "ms_per_step": ms_mf,
                   "ksp_iterations_per_step": its_mf,
                   "K_product_us": round(t_mf / max(n_mf, 1) * 1e3, 2),
                   "roofline": {"bound": "fp64 pipe / latency", "flop_per_product": int(2 * 24000 * fem.dom.getDeviceMesh().nel),
                                "flop_note": "~24 k FMA per element at ngl=4 (3 forward + 3 transposed sum-factorised "
                                             "passes per rule + the point updates, DESIGN.md section 3)",
                                "achieved_TFLOPs": round(2 * 24000 * fem.dom.getDeviceMesh().nel / (t_mf / max(n_mf, 1) * 1e-3) / 1e12, 2),
                                "peak_TFLOPs": FP64_PEAK["TFLOPs"],
                                "frac": round(2 * 24000 * fem.dom.getDeviceMesh().nel / (t_mf / max(n_mf, 1) * 1e-3) / 1e12
                                              / FP64_PEAK["TFLOPs"], 4) if FP64_PEAK["TFLOPs"] else None},
                   "K_product": "mf_apply_kernel (sum-factorised, fp64 pipe) + mf_gather_kernel; "
                                "vectors + 80 B of geometry per quadrature point per product"}
        if not args.no_e2e:
            matfree["e2e"] = run_e2e(args, fem, ctx, world, ndof, nvort, restore)
        K.setMatFree(False)

    out = {
        "metric": METRIC, "value": ndof / (ms_per_step * 1e-3), "unit": UNIT, "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": workload_name(args),
                   "dofs": int(ndof), "nodes": int(ndof // 3), "ksp": f"cg+{args.pc}", "ksp_rtol": args.rtol,
                   "ksp_options": f"-ksp_type cg -pc_type {args.pc}" +
                                  (f" -pc_asm_coarse {args.coarse}" if args.pc == "asm" else "") + f" -ksp_rtol {args.rtol:g}" +
                                  (f" -ksp_fischer_guess 1,{args.fischer}" if args.fischer > 0 else "") +
                                  "".join(" -" + kv.replace("=", " ") for kv in args.opt),
                   "dt": args.dt, "rhs_evals_per_step": evals / args.steps,
                   "ksp_iterations_per_step": its / args.steps, "ksp_solves_per_step": solves / args.steps,
                   "parallelism": "1 GPU" if world == 1 else
                   f"row partition over {world} GPUs ({'x'.join(str(p) for p in fem.dom._partition.pgrid)} element "
                   f"blocks), halo exchange + Krylov allreduces " +
                   ("fused into the kernels over peer memory (CUDA IPC / NVLink)" if ctx.p2p_status()[0]
                    else "over NCCL") + f", {ctx.collectives()} collectives so far on rank 0",
                   "l2": "inputs larger than L2 (K = %.2f GB per SpMV)" % (spmv_bytes / 1e9)},
        "roofline": roofline,
        "other_rooflines": other_rooflines,
        "fp64_peak": {"TFLOPs": FP64_PEAK["TFLOPs"], "how": "pyn_fp64_peak: 16 independent DFMA chains per thread, "
                      "8 CTAs of 256 threads per SM, best of 4 timed launches (2 flop per DFMA)"},
        "checksum": checksum,
        "kernel_time_ms_per_step": {"spmv_K_dot": ms_spmv / args.steps, "cg_update": ms_upd / args.steps,
                                    "cg_direction": ms_dir / args.steps, "spmv_other": ms_oth / args.steps},
        "assembly": assembly_report(asm, asm_kernels, fem, ndof, peak),
        "assembly_kernels_warm": assembly_warm(args, ctx, fem, peak) if not args.no_asm_warm else None,
        "gpu_launches": int(launches),
        "clocks": clocks,
    }
    if e2e:
        out["e2e"] = e2e
    if matfree:
        out["matfree"] = matfree
    if world == 1 and args.pc == "asm" and not args.no_alt:
        out["alt_pc_bjacobi"] = run_alt_pc(args, ctx, "bjacobi")
    return out


def run_alt_pc(args, ctx, pc):
    """the same K timed steps (after W warm-up steps from the initial condition) with another
    preconditioner -- node-block Jacobi, the round-1 default and PETSc's pbjacobi -- for comparison"""
    fem = build_problem(args, pc=pc)
    fem.ts.setSolution(fem.vort)
    for _ in range(args.warmup):
        fem.ts.step()
    ctx.sync()
    its0 = fem.kspIterations
    ctx.timer_start()
    for _ in range(args.steps):
        fem.ts.step()
    ms = ctx.timer_stop() / args.steps
    ndof = fem.mat.K.getSize()[0]
    res = {"ksp_options": f"-ksp_type cg -pc_type {pc} -ksp_rtol {args.rtol:g}" +
                          (f" -ksp_fischer_guess 1,{args.fischer}" if args.fischer > 0 else ""),
           "ms_per_step": ms, "value": ndof / (ms * 1e-3), "unit": UNIT,
           "ksp_iterations_per_step": (fem.kspIterations - its0) / args.steps}
    fem.mat.K.destroy()
    return res


def assembly_kernels(ctx, dm, mode, reps=3):
    """Re-assemble on an existing DeviceMesh `reps` times and return the kernel times of the LAST
    pass (warm: the first launch of a kernel also pays its module load between the two events)."""
    out = {}
    for _ in range(reps):
        ctx.sync()
        ctx.prof_enable(True)
        mats = dm.assemble_kle(mode)
        ops = dm.assemble_ops(compact=True)
        ctx.sync()
        out = {k: ctx.prof_get(k) for k in ("elem_kle", "numeric")}
        ctx.prof_enable(False)
        values = sum(m.nblocks() * m.getBlockSizes()[0] * m.getBlockSizes()[1] * 8
                     for m in list(mats.values()) + list(ops[:3]))
        for m in list(mats.values()) + list(ops[:3]):
            m.destroy()
        del mats, ops
    out["value_bytes"] = int(values)
    out["classes"] = dm.congruence()[1]
    return out


FP64_PEAK = {"TFLOPs": None}


def kernels_entry(k, peak, nel, ngl):
    n_el, ms_el = k["elem_kle"]
    n_nu, ms_nu = k["numeric"]
    # element kernel: 16 pair moments, 7 / 15 fp64 instructions per full / reduced point and pair
    # (DESIGN.md section 3), counted as 2 flop each, over the classes actually evaluated
    nn = ngl ** 3
    flop = 2.0 * nn * nn * (7 * ngl ** 3 + 15 * (ngl - 1) ** 3) * k["classes"]
    return {"congruence_classes": int(k["classes"]), "elements": int(nel),
            "elem_kle_ms": round(ms_el, 4), "elem_kle_TFLOPs": round(flop / (ms_el * 1e-3) / 1e12, 2) if ms_el else None,
            "elem_kle_frac_of_measured_fp64_peak": round(flop / (ms_el * 1e-3) / 1e12 / FP64_PEAK["TFLOPs"], 4)
            if ms_el and FP64_PEAK["TFLOPs"] else None,
            "numeric_ms": round(ms_nu, 4), "numeric_launches": n_nu, "numeric_value_bytes": k["value_bytes"],
            "numeric_GBs": round(k["value_bytes"] / (ms_nu * 1e-3) / 1e9, 1) if ms_nu else None,
            "numeric_frac_of_peak": round(k["value_bytes"] / (ms_nu * 1e-3) / 1e9 / peak, 4) if ms_nu else None}


def assembly_warm(args, ctx, fem, peak):
    """Warm kernel times of the assembly, (a) on the bench mesh as the case builds it (uniform
    box: ONE congruence class, the numeric pass is a pure HBM writer) and (b) on the perturbed
    twin of the same mesh (every interior corner displaced by 0.15 h, no two
    cells congruent: the graded input of the element kernel)."""
    from pynama_b200.device_mesh import DeviceMesh
    from pynama_b200.domain.box_mesh import BoxMesh
    dm = fem.dom.getDeviceMesh()
    mode = "NS" if getattr(fem.mat, "Kfs", None) is not None else "FS"
    res = {"note": "third of three back-to-back assemblies on the same mesh tables; CUDA events around the "
                   "element kernel and the numeric (COO -> node-block BSR) kernels; bytes = 8 B per stored "
                   "value written once (SURVEY.md 8d-D4)",
           "uniform": kernels_entry(assembly_kernels(ctx, dm, mode), peak, dm.nel, args.ngl)}
    if int(os.environ.get("WORLD_SIZE", "1")) == 1:
        m = fem.dom._mesh
        pm = BoxMesh(m.nelem, m.lower, m.upper, m.ngl, perturb=0.15)
        pdm = DeviceMesh(m.dim, m.ngl, pm.corner_xyz, pm.elem_nodes, pm.nnodes)
        pdm.set_dof_class(fem.dom.getDofClasses())
        res["perturbed"] = kernels_entry(assembly_kernels(ctx, pdm, mode), peak, pdm.nel, args.ngl)
        pdm.destroy()
    return res


def assembly_report(asm, kern, fem, ndof, peak):
    """phase_ms: CUDA-event time of whole phases (first call in a fresh process: includes
    cudaMalloc of the matrices); kernels: event time of the element kernel (fp64-bound) and of the
    numeric COO->BSR kernels alone, with the numeric phase's algorithmic bytes = 8 B per stored
    value written once."""
    mats = [fem.mat.K, fem.mat.Krhs, fem.mat.Rw, fem.operator.Curl, fem.operator.SrT, fem.operator.DivSrT]
    stored = sum(m.bytes()[0] for m in mats)
    values = sum(m.nblocks() * m.getBlockSizes()[0] * m.getBlockSizes()[1] * 8 for m in mats)
    ref_nnz = sum(m.nnz()[0] for m in mats)
    total_ms = sum(asm.values())
    n_el, ms_el = kern["elem_kle"]
    n_nu, ms_nu = kern["numeric"]
    return {"phase_ms": {k: round(v, 3) for k, v in asm.items()}, "total_ms": round(total_ms, 3),
            "dof_per_s": ndof / (total_ms * 1e-3) if total_ms else None,
            "stored_bytes": int(stored), "reference_pattern_nnz": int(ref_nnz),
            "kernels": {"elem_kle_ms": round(ms_el, 3), "numeric_ms": round(ms_nu, 3), "numeric_launches": n_nu,
                        "numeric_value_bytes": int(values),
                        "numeric_GBs": round(values / (ms_nu * 1e-3) / 1e9, 1) if ms_nu else None,
                        "numeric_frac_of_peak": round(values / (ms_nu * 1e-3) / 1e9 / peak, 4) if ms_nu else None},
            "host_setup_wall_s": round(fem.setup_wall_s, 3)}


def run_e2e(args, fem, ctx, world, ndof, nvort, restore=None):
    """public-API time step with HOST-resident fields: upload vorticity+velocity from pinned memory,
    ts.step(), read both back."""
    vel = fem.solverKLE.getSolution()
    h_vort, h_vel = ctx.pinned_empty(fem.vort.getLocalSize()), ctx.pinned_empty(vel.getLocalSize())
    fem.vort.getArray(out=h_vort)
    vel.getArray(out=h_vel)
    bcb = bc_bytes_per_rhs(fem)
    evals0 = fem.ts.rhs_evals

    def step():
        fem.vort.setArray(h_vort)
        vel.setArray(h_vel)
        fem.ts.step()
        fem.vort.getArray(out=h_vort)
        vel.getArray(out=h_vel)

    if restore is not None:            # time the same K steps as the device-resident regions
        restore()
        fem.vort.getArray(out=h_vort)
        vel.getArray(out=h_vel)
    else:
        step()
    ctx.sync()
    barrier(world)
    evals0 = fem.ts.rhs_evals
    ctx.timer_start()
    for _ in range(args.steps):
        step()
    ms = ctx.timer_stop()
    ms = max_over_ranks(ms, world)
    evals = (fem.ts.rhs_evals - evals0) / args.steps
    nloc = h_vort.size + h_vel.size
    return {"value": ndof / (ms / args.steps * 1e-3), "unit": UNIT, "ms_per_step": ms / args.steps,
            "h2d_bytes_per_step": int(nloc * 8 + evals * bcb),
            "d2h_bytes_per_step": int(nloc * 8), "bytes_are": "per rank (rank 0)" if world > 1 else "total"}


def read_traffic():
    """dram bytes per launch of the dominant kernel from the committed ncu capture, if any"""
    p = os.path.join(ROOT, "profiles", "traffic.json")
    try:
        with open(p) as f:
            return json.load(f).get("spmv_K_dram_bytes_per_launch")
    except Exception:
        return None


# ---------------------------------------------------------------------------------------------
# CPU: the oracle (a port of the reference's algorithm; the reference itself needs petsc4py,
# which is not installed) timed on the host cores.  Checker code, used here only as the
# reported baseline / the --impl reference arm.
# ---------------------------------------------------------------------------------------------
CHECK_STEPS = 2          # the checksum is taken after this many steps from the initial condition


def field_checksum(vort, vel):
    """norms of the fields (numpy, whole vectors): compared between the GPU arm at every N and the CPU arm"""
    return {"vort_l2": float(np.sqrt(vort @ vort)), "vel_l2": float(np.sqrt(vel @ vel)),
            "vort_max": float(np.abs(vort).max())}


def cpu_pc(args):
    """the CPU arm's preconditioner: the GPU arm's when the C port has it, else node-block Jacobi (PETSc's
    pbjacobi) -- the Schwarz preconditioner of the GPU arm has no CPU port; same KSP type, tolerance,
    Fischer guess, mesh and time step either way"""
    return args.pc if args.pc in ("jacobi", "bjacobi") else "bjacobi"


def cpu_problem(args):
    """the CPU restatement on the SAME mesh as the GPU arm: oracle/boxmesh.py tables, oracle/sem.py cell
    matrices, oracle/ckle.c assembly + scalar-CSR products + CG on all host cores (oracle/cport.py).
    No import of the product package."""
    from oracle import boxmesh, cport
    t0 = time.perf_counter()
    mesh = boxmesh.box_mesh([args.nelem] * 3, [0.0] * 3, [1.0] * 3, args.ngl)
    prob = cport.CProblem(3, args.ngl, mesh.corner_xyz, mesh.elem_nodes, mesh.coords, mesh.border_nodes,
                          rho=0.5, mu=0.01, rtol=args.rtol, pc=cpu_pc(args), fischer=args.fischer)
    return prob, mesh.nnodes * 3, time.perf_counter() - t0, cport.threads()


def cpu_run(args, steps, warmup, budget_s):
    """W' <= warmup untimed and K' <= steps timed RK5BS steps of the full problem, cut short when the
    wall-clock budget is spent (what ran is what is reported).  Returns a dict."""
    prob, ndof, setup_s, cores = cpu_problem(args)
    t_all = time.perf_counter()
    t, k0, chk = 0.0, None, None
    done_w = 0
    for _ in range(warmup):
        if done_w >= 1 and time.perf_counter() - t_all > 0.35 * budget_s:
            break
        t, k0 = prob.step(t, args.dt, k0)
        done_w += 1
        if done_w == CHECK_STEPS:
            chk = field_checksum(prob.vort, prob.vel)
    its0 = prob.iterations
    t0 = time.perf_counter()
    done = 0
    for _ in range(steps):
        if done >= 1 and time.perf_counter() - t_all > budget_s:
            break
        t, k0 = prob.step(t, args.dt, k0)
        done += 1
        if done_w + done == CHECK_STEPS:
            chk = field_checksum(prob.vort, prob.vel)
    wall = time.perf_counter() - t0
    ms = wall / done * 1e3
    its = (prob.iterations - its0) / done
    sample = (f"{done} timed RK5BS step(s) after {done_w} warm-up step(s) of the SAME problem ({args.nelem}^3 elements, "
              f"ngl={args.ngl}, {ndof} dofs): cell matrices from the numpy oracle, scatter-add assembly, scalar-CSR "
              f"SpMV and CG + {cpu_pc(args)} (rtol {args.rtol:g}" +
              (f", Fischer guess 1,{args.fischer}" if args.fischer > 0 else "") +
              f") in C + OpenMP on {cores} threads; {its:.0f} CG its/step, {wall:.1f} s timed, "
              f"{setup_s:.1f} s set-up + assembly")
    return {"value": ndof / (ms * 1e-3), "ms": ms, "sample": sample, "cores": cores, "steps": done,
            "warmup": done_w, "its": its, "ndof": ndof, "setup_s": setup_s, "checksum": chk,
            "K_nnz": int(prob.K.nnz)}


def run_reference(args, rank, world):
    """--impl reference: the reference's own implementation needs petsc4py + mpirun, which neither
    machine has, so this arm times the CPU restatement of the path on all host
    cores, on the same mesh, options and metric as the GPU arm."""
    if rank != 0:
        return None
    r = cpu_run(args, args.steps, args.warmup, budget_s=float(os.environ.get("PYN_CPU_BUDGET_S", "200")))
    return {"impl": "reference", "metric": METRIC, "value": r["value"], "unit": UNIT, "n_gpus": world,
            "steps": r["steps"], "warmup": r["warmup"], "steps_requested": args.steps,
            "warmup_requested": args.warmup, "ms_per_step": r["ms"], "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": workload_name(args), "dofs": int(r["ndof"]), "ksp": f"cg+{cpu_pc(args)}",
                       "ksp_rtol": args.rtol, "dt": args.dt, "ksp_iterations_per_step": r["its"],
                       "K_scalar_nnz": r["K_nnz"],
                       "note": "petsc4py/PETSc/mpirun are not installed: the reference's own modules cannot be "
                               "imported; this arm times the CPU restatement (oracle/) of the same path. steps / "
                               "warmup are what actually ran inside the wall-clock budget"},
            "checksum": dict(r["checksum"] or {}, after_steps=CHECK_STEPS),
            "cpu_baseline": {"value": r["value"], "unit": UNIT, "cores": r["cores"], "kind": "port",
                             "sample": r["sample"]},
            "e2e": {"value": r["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}


# ---------------------------------------------------------------------------------------------
def init_dist():
    """rank / world size from the torchrun environment; the rendezvous itself (gloo on 127.0.0.1,
    used once to ship the NCCL id) happens in pynama_b200.petsc.Context"""
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    # rank 0 prints ONE JSON line on stdout: keep NCCL's version banner (NCCL_DEBUG=VERSION) out of it
    if os.environ.get("NCCL_DEBUG", "").upper() in ("", "VERSION"):
        os.environ["NCCL_DEBUG"] = "WARN"
    return rank, world


def barrier(world):
    if world > 1:
        from pynama_b200.petsc import Context
        ctx = Context.get()
        ctx.sync()
        ctx.allreduce([0.0], "sum")          # NCCL allreduce + host read-back: all ranks arrived
        ctx.sync()


def max_over_ranks(x, world):
    if world > 1:
        from pynama_b200.petsc import Context
        return float(Context.get().allreduce([x], "max")[0])
    return x


def main():
    args = parse_args()
    rank, world = init_dist()
    if args.impl == "reference":
        out = run_reference(args, rank, world)
    else:
        out = run_b200(args, rank, world)
        if rank == 0 and world == 1 and not args.no_cpu_baseline:
            r = cpu_run(args, 1, 1, budget_s=120.0)            # bounded sample: 1 warm-up + 1 timed step
            out["cpu_baseline"] = {"value": r["value"], "unit": UNIT, "cores": r["cores"], "kind": "port",
                                   "sample": r["sample"]}
            if r["checksum"] and out.get("checksum"):
                g = out["checksum"]
                out["cpu_baseline"]["checksum"] = r["checksum"]
                out["cpu_baseline"]["checksum_rel_diff_vs_gpu"] = max(
                    abs(g[k] - r["checksum"][k]) / abs(r["checksum"][k]) for k in r["checksum"])
    if rank == 0 and out is not None:
        print(json.dumps(out), flush=True)
    if world > 1:
        try:
            import torch.distributed as dist
            if dist.is_initialized():
                dist.barrier()
                dist.destroy_process_group()
        except Exception:
            pass


if __name__ == "__main__":
    main()
