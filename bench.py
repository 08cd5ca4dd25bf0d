#!/usr/bin/env python
"""Benchmark of the step-35a time-step path on B200.

    python bench.py --gpus N --steps K --warmup W [--impl reference] [--mesh M --dim D --refines R]

One "step" is one pass of the loop body of NavierStokesProjection::run (step-35a.cc:1190-1239): diffusion step
(per-component ILU(0)-GMRES), velocity extrapolation, advection assembly, projection step (ILU(0)-CG) and
pressure update, on the 3D cylinder mesh (Q2/Q1, parameter-file.prm values with pressure_fe_degree = 1).
Rank 0 prints ONE JSON line.  `--impl reference` times the CPU oracle (the reference itself needs deal.II and
cannot be built here) on the host cores with the same workload, metric and unit.
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
sys.path.insert(0, os.path.join(ROOT, "tests"))
import __graft_entry__ as ge  # noqa: E402

REYNOLDS = {"cylinderextruded": 100.0, "cylinder3d": 20.0, "nsbench2": 100.0}

def ncu_traffic(workload, kernel):
    """dram__bytes_read.sum + dram__bytes_write.sum per application of `kernel` on `workload`, from the summaries of the
    committed `ncu --set full` captures (profiles/ncu_traffic.json, written by tools/ncu_summary.py); None if that
    kernel was not captured on that workload."""
    path = os.path.join(ROOT, "profiles", "ncu_traffic.json")
    if not os.path.exists(path):
        return None
    with open(path) as f:
        table = json.load(f)
    entry = table.get(workload, {}).get(kernel)
    return entry["dram_bytes"] if entry else None


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--mesh", default="cylinderextruded")
    ap.add_argument("--dim", type=int, default=3)
    ap.add_argument("--refines", type=int, default=1)
    ap.add_argument("--cpu-seconds", type=float, default=15.0, help="budget of the cpu_baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-kernels", action="store_true", help="skip the per-kernel roofline pass")
    ap.add_argument("--no-parity", action="store_true", help="skip the field comparison with the CPU oracle")
    ap.add_argument("--parity-steps", type=int, default=0, help="steps compared with the oracle (default: 3 on one GPU, 2 on several)")
    ap.add_argument("--ilu-blocks", type=int, default=1,
                    help="GATED ALTERNATIVE (not the reference's preconditioner): ILU(0) of k index-range blocks")
    ap.add_argument("--off-diagonals", type=int, default=0,
                    help="off_diagonals of parameter-file.prm (0 there and here; the reference's Readme reports 5 as faster)")
    ap.add_argument("--deadline", type=int, default=600, help="abort the whole run after this many seconds")
    return ap.parse_args()


def load_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks and throttle reasons while the timed region runs."""

    QUERY = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")

    def __init__(self, device_index):
        self.device_index = device_index
        self.lines = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.device_index), "--query-gpu=" + self.QUERY,
                                          "--format=csv,noheader,nounits", "-lms", "200"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        sm, smax, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in self.lines:
            f = [x.strip() for x in line.split(",")]
            if len(f) < 8:
                continue
            try:
                sm.append(float(f[1]))
                smax.append(float(f[2]))
            except ValueError:
                continue
            for name, v in zip(names, f[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(smax) if smax else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def workload_name(args):
    return "%s r%d %dD Q2/Q1 Re=%g" % (args.mesh, args.refines, args.dim, REYNOLDS.get(args.mesh, 100.0))


def build_disc(pkg, args):
    params = pkg.default_params(reynolds=REYNOLDS.get(args.mesh, 100.0), final_time=1.0, dt=5e-4, vel_eps=1e-8,
                                vel_off_diagonals=args.off_diagonals, vel_diag_strength=0.0, vel_update_prec=100, vel_krylov_size=30,
                                vel_max_iterations=1000, n_global_refines=args.refines, pressure_degree=1, rotational=1)
    mesh_file = os.path.join(ROOT, "meshes", args.mesh + ".s35m")
    return pkg.Disc(mesh_file, args.dim, params), params


def preconditioner_label(args, world):
    if args.ilu_blocks > 1:
        return "GATED ALTERNATIVE: ILU(0) of %d index-range blocks" % args.ilu_blocks
    if world > 1:
        return ("block-local ILU(0) per rank (couplings to ghost DoFs dropped; NOT the reference's global SparseILU: "
                "iteration counts differ from the single-GPU run)")
    if args.off_diagonals > 0:
        return "ILU(0) on the pattern widened by off_diagonals = %d, as the reference's SparseILU with that setting" % args.off_diagonals
    return "ILU(0) as the reference's SparseILU"


def config_dict(args, disc, world):
    """the same for both arms: it describes the workload, not who runs it"""
    dim = disc.dim
    return {"workload": workload_name(args), "dof": dim * disc.nv + disc.np_, "n_vel_per_component": disc.nv, "n_pres": disc.np_,
            "cells": disc.n_cells, "preconditioner": preconditioner_label(args, world),
            # every S35_* switch that was set: a line measured with a non-default path says so
            "switches": {k: v for k, v in sorted(os.environ.items()) if k.startswith("S35_")},
            "parallelism": ("mesh partition x%d (coordinate bisection, NCCL ghost exchange + all-reduce, "
                            "block-local ILU(0))" % world) if world > 1 else "single GPU",
            "l2": "working set (%.0f MB of matrices) exceeds L2; kernel benchmarks flush L2 between launches"
                  % ((disc.info.nnz_vel_constrained * (dim + 1) * 8 + disc.info.nnz_vel_constrained * 4) / 1e6)}


def field_differences(pkg, disc, reynolds, u_a, p_a, u_b, p_b):
    """largest and l2 difference of two states relative to the second one's norms, and of their drag / lift"""
    out = {}
    du = max(np.abs(a - b).max() for a, b in zip(u_a, u_b))
    out["u_rel_max"] = float(du / max(max(np.abs(b).max() for b in u_b), 1e-300))
    out["u_rel_l2"] = float(np.sqrt(sum(((a - b) ** 2).sum() for a, b in zip(u_a, u_b)) / max(sum((b ** 2).sum() for b in u_b), 1e-300)))
    out["p_rel_max"] = float(np.abs(p_a - p_b).max() / max(np.abs(p_b).max(), 1e-300))
    out["p_rel_l2"] = float(np.sqrt(((p_a - p_b) ** 2).sum() / max((p_b ** 2).sum(), 1e-300)))
    cd_a = disc.drag_lift(reynolds, np.concatenate(u_a), p_a)
    cd_b = disc.drag_lift(reynolds, np.concatenate(u_b), p_b)
    out["cd_cl"] = [float(x) for x in cd_b]
    out["cd_cl_abs"] = float(np.abs(cd_a - cd_b).max())
    return out


def iteration_counts(c, dim):
    return {"gmres": [c.solver_stats(d).iterations for d in range(dim)], "cg": c.solver_stats(-1).iterations}


class CpuArm:
    """The oracle (-O3 -march=native, OpenMP over cells / components / rows like the reference's TBB use) on the host
    cores, advanced from the initial state like the device run."""

    def __init__(self, pkg, disc):
        import oracle_binding as ob

        self.pkg, self.disc = pkg, disc
        self.cores = os.cpu_count() or 1
        self.orc = ob.Oracle(disc.problem, fast=True, threads=self.cores)
        self.orc.set_field(pkg.PRES_N, disc.initial_pressure)
        self.orc.set_field(pkg.PRES_N_MINUS_1, disc.initial_pressure)
        self.n = 2

    def advance(self, steps):
        for _ in range(steps):
            self.orc.advance(self.n, 1)
            self.n += 1

    def state(self):
        pkg = self.pkg
        return [self.orc.get_field(pkg.U_N, d) for d in range(self.disc.dim)], self.orc.get_field(pkg.PRES_N)

    def timed(self, budget_seconds, steps=None):
        """steps/s over a bounded sample of steady steps (whatever was advanced before is the warm-up)"""
        sec0 = {s: self.orc.section_seconds(s) for s in self.pkg.SECTIONS}
        done, t0 = 0, time.perf_counter()
        while True:
            self.advance(1)
            done += 1
            elapsed = time.perf_counter() - t0
            if steps is not None:
                if done >= steps:
                    break
            elif elapsed >= budget_seconds or done >= 50:
                break
        sections = {s: self.orc.section_seconds(s) - sec0[s] for s in self.pkg.SECTIONS}
        return done / elapsed, done, sections, iteration_counts(self.orc, self.disc.dim)

    def close(self):
        self.orc.close()


def main():
    args = parse_args()
    # stdout carries exactly one JSON line: everything else that lands on file descriptor 1 (the NCCL version
    # banner, library chatter) is sent to stderr, and the line is written to a private copy of the real stdout
    real_stdout = os.fdopen(os.dup(1), "w")
    sys.stdout.flush()
    os.dup2(2, 1)

    def emit(line):
        real_stdout.write(json.dumps(line) + "\n")
        real_stdout.flush()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    # NCCL prints its version banner on stdout when NCCL_DEBUG=VERSION: keep stdout to the one JSON line
    if os.environ.get("NCCL_DEBUG", "VERSION").upper() == "VERSION":
        os.environ["NCCL_DEBUG"] = "WARN"
    # the collectives of this path are 8..240-byte reductions and ghost layers of a few thousand values:
    # in-switch reduction (NVLS) buys nothing there and costs set-up time per communicator
    os.environ.setdefault("NCCL_NVLS_ENABLE", "0")
    # a dead-locked collective must not hold the GPUs for ever: give up after --deadline seconds
    def give_up():
        sys.stderr.write("bench.py: deadline of %d s exceeded on rank %d, aborting\n" % (args.deadline, rank))
        sys.stderr.flush()
        os._exit(3)
    watchdog = threading.Timer(args.deadline, give_up)
    watchdog.daemon = True
    watchdog.start()
    # torchrun pins OMP_NUM_THREADS to 1; the host set-up (OpenMP) may use this rank's share of the cores
    os.environ["OMP_NUM_THREADS"] = str(max(1, (os.cpu_count() or 1) // max(1, world)))
    pkg = ge.load_package()

    if args.impl == "reference":
        if rank != 0:
            return 0
        disc, _ = build_disc(pkg, args)
        dof = disc.dim * disc.nv + disc.np_
        arm = CpuArm(pkg, disc)
        arm.advance(args.warmup)
        rate, done, sections, its = arm.timed(args.cpu_seconds, steps=args.steps)
        cores = arm.cores
        arm.close()
        line = {
            "impl": "reference", "metric": "time_steps_per_s", "value": rate, "unit": "steps/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 / rate, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": config_dict(args, disc, args.gpus),
            "dof_steps_per_s": rate * dof,
            "cpu_baseline": {"value": rate, "unit": "steps/s", "cores": cores, "kind": "port",
                             "sample": "%d warm-up + %d timed steps of the same workload; oracle/oracle.cpp built -O3 "
                                       "-march=native with OpenMP (the reference needs deal.II: unavailable on this box)"
                                       % (args.warmup, done)},
            "e2e": {"value": rate, "unit": "steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "sections_s": sections,
            # what this process mapped of the repository's own native code: the host-only set-up library and the oracle
            "native_libraries": sorted({os.path.basename(ln.split()[-1]) for ln in open("/proc/self/maps")
                                        if ROOT in ln and ".so" in ln}),
            "iterations": its,
        }
        emit(line)
        return 0

    import torch
    import torch.distributed as dist

    if pkg.device_count() < 1:
        raise RuntimeError("bench.py: no CUDA device; the device path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    pkg.check(pkg.lib().s35_set_device(local_rank))
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    disc, params = build_disc(pkg, args)
    dim = disc.dim
    dof = dim * disc.nv + disc.np_
    part = None
    if world > 1:
        # the mesh is partitioned over the ranks; the library's own NCCL communicator carries the ghost exchange and
        # the Krylov reductions (torch.distributed only hands out the communicator id and synchronises the timing)
        uid = torch.zeros(128, dtype=torch.uint8, device="cuda")
        if rank == 0:
            uid = torch.tensor(list(pkg.comm_unique_id()), dtype=torch.uint8, device="cuda")
        dist.broadcast(uid, 0)
        pkg.comm_init(rank, world, bytes(uid.cpu().tolist()))
        part = pkg.Part(disc, world, rank)
    problem = part.problem if part else disc.problem
    if args.ilu_blocks > 1:
        problem = pkg.with_ilu_blocks(problem, args.ilu_blocks)
    initial_pressure = part.initial_pressure if part else disc.initial_pressure
    nv, np_ = (part.nv, part.np_) if part else (disc.nv, disc.np_)
    t0 = time.perf_counter()
    ctx = pkg.Ctx(problem)
    ctx_seconds = time.perf_counter() - t0
    ctx.set_field(pkg.PRES_N, initial_pressure)
    ctx.set_field(pkg.PRES_N_MINUS_1, initial_pressure)

    # ---- parity on the benchmarked workload: the first steps from the initial state on the device and in the CPU oracle
    # (rank 0; the oracle always solves the unpartitioned problem with the reference's global ILU(0)), fields compared by
    # global DoF index.  These steps double as warm-up for both arms.
    n = 2
    parity, cpu = None, None
    want_cpu = rank == 0 and not args.no_cpu_baseline
    if not args.no_parity:
        p_steps = args.parity_steps or (3 if world == 1 else 2)
        ctx.advance(n, p_steps)
        n += p_steps
        its_gpu = iteration_counts(ctx, dim)
        u_own = [ctx.get_field(pkg.U_N, d) for d in range(dim)]
        p_own = ctx.get_field(pkg.PRES_N)
        if world > 1:
            # every rank contributes the entries it owns (the first nv_owned / np_owned of its local vectors)
            gathered = [None] * world if rank == 0 else None
            mine = {"vel": np.array(part.vel_global[:part.nv_owned]), "pres": np.array(part.pres_global[:part.np_owned]),
                    "u": [u[:part.nv_owned].copy() for u in u_own], "p": p_own[:part.np_owned].copy()}
            dist.gather_object(mine, gathered, dst=0)
            if rank == 0:
                u_dev = [np.zeros(disc.nv) for _ in range(dim)]
                p_dev = np.zeros(disc.np_)
                for g in gathered:
                    for d in range(dim):
                        u_dev[d][g["vel"]] = g["u"][d]
                    p_dev[g["pres"]] = g["p"]
        else:
            u_dev, p_dev = u_own, p_own
        if rank == 0:
            cpu = CpuArm(pkg, disc)
            cpu.advance(p_steps)
            u_cpu, p_cpu = cpu.state()
            parity = field_differences(pkg, disc, REYNOLDS.get(args.mesh, 100.0), u_dev, p_dev, u_cpu, p_cpu)
            parity.update({"steps": p_steps, "its_gpu": its_gpu, "its_cpu": iteration_counts(cpu.orc, dim),
                           "against": "oracle/oracle.cpp from the same initial state (global ILU(0), one process)",
                           "note": "the solvers stop at eps * ||rhs||: fields agree to about eps x condition number "
                                   "(BASELINE.md section 4), drag/lift to 1e-6"})
            if not want_cpu:
                cpu.close()
                cpu = None
        barrier()

    # ---- device-resident run: W warm-up steps, then exactly K timed steps
    if args.warmup:
        ctx.advance(n, args.warmup)
        n += args.warmup
    sampler = ClockSampler(local_rank)
    barrier()
    sampler.start()
    launches0 = pkg.launch_count()
    sec0 = {s: ctx.section_ms(s)[0] for s in pkg.SECTIONS}
    t0 = time.perf_counter()
    dev_ms = ctx.advance(n, args.steps)
    barrier()
    wall_s = time.perf_counter() - t0
    launches = pkg.launch_count() - launches0
    clocks = sampler.stop()
    n += args.steps
    its = iteration_counts(ctx, dim)
    sections = {s: (ctx.section_ms(s)[0] - sec0[s]) / args.steps for s in pkg.SECTIONS}
    elapsed = max(dev_ms * 1e-3, 0.0)
    t = torch.tensor([elapsed, wall_s], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    elapsed, wall_s = float(t[0]), float(t[1])
    # one partitioned problem advanced by all ranks together: whole-job steps/s
    value = args.steps / elapsed

    # ---- end to end through the C ABI with host buffers: the state a host-side caller owns (the checkpoint set of
    # step-35a.cc:865-875) goes to the device from pinned memory every step and comes back after it
    e2e = None
    fields_v = (pkg.U_N, pkg.U_N_MINUS_1)
    fields_p = (pkg.PRES_N, pkg.PRES_N_MINUS_1, pkg.PHI_N, pkg.PHI_N_MINUS_1)
    host = {}
    for f in fields_v:
        for d in range(dim):
            buf = torch.empty(nv, dtype=torch.float64, pin_memory=True).numpy()
            buf[:] = ctx.get_field(f, d)
            host[(f, d)] = buf
    for f in fields_p:
        buf = torch.empty(np_, dtype=torch.float64, pin_memory=True).numpy()
        buf[:] = ctx.get_field(f)
        host[(f, 0)] = buf
    lib = pkg.lib()

    def e2e_step(step_n):
        for (f, d), buf in host.items():
            pkg.check(lib.s35_ctx_set_field(ctx.handle, f, d, pkg.dptr(buf)))
        ctx.advance(step_n, 1, start_n=-1, start_offset_n=-1)
        for (f, d), buf in host.items():
            pkg.check(lib.s35_ctx_get_field(ctx.handle, f, d, pkg.dptr(buf)))

    e2e_steps = max(1, min(args.steps, 5))
    e2e_step(n)
    n += 1
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        e2e_step(n)
        n += 1
    barrier()
    e2e_s = time.perf_counter() - t0
    t = torch.tensor([e2e_s], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    state_bytes = 8 * (len(fields_v) * dim * nv + len(fields_p) * np_)
    sb = torch.tensor([float(state_bytes)], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(sb)
    state_bytes = int(sb[0])
    e2e = {"value": e2e_steps / float(t[0]), "unit": "steps/s", "h2d_bytes_per_step": state_bytes,
           "d2h_bytes_per_step": state_bytes, "steps": e2e_steps}

    # ---- per-kernel rooflines (CUDA events on the launch stream, L2 flushed between launches)
    peak, peak_kind = load_peaks()
    kernels = {}
    fp64_peak_tflops = None
    if not args.no_kernels:
        names = {"spmv_vel": pkg.BENCH_SPMV_VEL, "spmv_pres": pkg.BENCH_SPMV_PRES, "spmv_grad": pkg.BENCH_SPMV_GRAD,
                 "spmv_grad_T": pkg.BENCH_SPMV_GRAD_T, "assemble_advection": pkg.BENCH_ASSEMBLE,
                 "ilu_apply_vel": pkg.BENCH_ILU_APPLY_VEL, "ilu_apply_pres": pkg.BENCH_ILU_APPLY_PRES,
                 "dot": pkg.BENCH_DOT, "add_and_dot": pkg.BENCH_ADD_AND_DOT}
        for name, kid in names.items():
            ms, nbytes, nl = ctx.bench_kernel(kid, warmup=3, reps=10, flush_l2=True)
            gbs = nbytes / (ms * 1e-3) / 1e9
            kernels[name] = {"ms": ms, "algorithmic_bytes": nbytes, "GB/s": gbs, "frac": gbs / peak, "launches": nl}
        # the assembly is FP64-FMA work: 2 (dim + 1) dpc^3 flop per cell against the FMA rate
        # measured here, now
        ms, flops, _ = ctx.bench_kernel(pkg.BENCH_FP64_FMA, warmup=2, reps=5, flush_l2=False)
        fp64_peak_tflops = flops / (ms * 1e-3) / 1e12
        dpc = 3 ** disc.dim
        asm_flops = 2.0 * (disc.dim + 1) * dpc ** 3 * disc.n_cells
        a = kernels["assemble_advection"]
        a.update({"flops": asm_flops, "TFLOP/s": asm_flops / (a["ms"] * 1e-3) / 1e12,
                  "frac_fp64": asm_flops / (a["ms"] * 1e-3) / 1e12 / fp64_peak_tflops})
    # the kernel that dominates the step: the ILU(0) substitution sweeps (SpMV is reported beside it)
    roofline = None
    if kernels:
        per_step = {
            # the dim component solves run concurrently (a stream each): the longest one counts
            "ilu_apply_vel": kernels["ilu_apply_vel"]["ms"] * (max(its["gmres"]) + 1),
            "ilu_apply_pres": kernels["ilu_apply_pres"]["ms"] * (its["cg"] + 1),
            "spmv_vel": kernels["spmv_vel"]["ms"] * (max(its["gmres"]) + 1),
            "assemble_advection": kernels["assemble_advection"]["ms"],
        }
        dominant = max(per_step, key=per_step.get)
        k = kernels[dominant]
        roofline = {"kernel": dominant, "bound": "hbm", "achieved": k["GB/s"], "peak": peak, "unit": "GB/s",
                    "frac": k["frac"], "traffic": ncu_traffic(workload_name(args), dominant) if world == 1 else None,
                    "algorithmic_bytes": k["algorithmic_bytes"], "ms": k["ms"], "launches": k["launches"],
                    "peak_source": peak_kind,
                    "share_of_step": per_step[dominant] / (1e3 * elapsed / args.steps),
                    "note": "critical-path bound (dependency levels x hand-off latency), not bandwidth bound: see DESIGN.md section 3"
                    if dominant.startswith("ilu") else ""}

    cpu_baseline = None
    if want_cpu and world == 1:
        if cpu is None:
            cpu = CpuArm(pkg, disc)
        # steady steps only, like the device arm: whatever the parity run did not cover of the warm-up comes first
        cpu.advance(max(0, args.warmup - (cpu.n - 2)))
        warm = cpu.n - 2
        rate, done, cpu_sections, cpu_its = cpu.timed(args.cpu_seconds)
        cpu_baseline = {"value": rate, "unit": "steps/s", "cores": cpu.cores, "kind": "port",
                        "sample": "%d timed time steps of the same workload after %d warm-up steps from the same initial "
                                  "state (oracle/oracle.cpp, -O3 -march=native, OpenMP)" % (done, warm),
                        "sections_s_per_step": {k: v / done for k, v in cpu_sections.items()}, "iterations": cpu_its}
    if cpu is not None:
        cpu.close()

    if rank == 0:
        line = {
            "metric": "time_steps_per_s", "value": value, "unit": "steps/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * elapsed / args.steps, "higher_is_better": True,
            # one mesh of fixed size is partitioned over the ranks: the total work does not grow with N
            "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": config_dict(args, disc, world),
            "dof_steps_per_s": value * dof, "wall_ms_per_step": 1e3 * wall_s / args.steps,
            "sections_ms_per_step": sections, "gpu_launches": launches, "clocks": clocks, "e2e": e2e, "roofline": roofline,
            "kernels": kernels, "fp64_peak_tflops": fp64_peak_tflops, "cpu_baseline": cpu_baseline,
            "setup_s": {"host_discretisation": disc.info.setup_seconds, "device_context": ctx_seconds},
            # last, so that a truncated tail of the line keeps them
            "iterations": its, "parity": parity,
        }
        emit(line)
    ctx.close()
    if world > 1:
        pkg.comm_finalize()
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
