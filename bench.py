#!/usr/bin/env python
"""bench.py -- headline benchmark: SOR MLUPS (and time steps/s) of the lid-driven-cavity time step.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--size 8192]

Workload (BASELINE.json configs[2], the one the metric is quoted on; SURVEY 8d "C3"): synthetic
lid-driven cavity, 8192 x 8192 cells, FP64, Re=1000, a=b=1, tau=1, omega=1.7, eps=1e-4, max_it=500.
Deterministic: all fields start at zero, lid u=1 -- no RNG.  One "step" = one whole time step
(dt reduction, boundary conditions, F/G, RHS, red-black SOR with the residual test every iteration
-- 500 iterations, it cannot converge at this size -- velocity update).  1 LUP = one interior cell
receiving its red-or-black update once; MLUPS counts SOR iterations actually performed x cells.

Timed region: K steps bracketed by barrier + synchronize, CUDA events on the stream the kernels
are launched on, max over ranks.  L2: one pressure grid is 537 MB >> 126 MB L2, every iteration
streams 1.6 GB, so inputs are larger than L2 (no flush needed; stated in config.l2).

Also reported: e2e (same steps through the host-grid call nsb_timestep with HOST buffers, pinned,
H2D/D2H inside the timed region), roofline of the dominant kernel (sor_stream_kernel, 24 B per
LUP against the measured HBM copy bandwidth), cpu_baseline (the unmodified reference serial SOR,
oracle/_ref, 1 core, bounded sample).  --impl reference times only that CPU reference.
"""
import argparse
import ctypes as C
import importlib
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402


def pkg():
    return importlib.import_module("navierstokes-parallel_b200")


def workload_params(m_or_ob, size, max_it):
    return m_or_ob.make_params(problem=1, f=1.0, i_max=size, j_max=size, a=1.0, b=1.0, T=1.0e9, Re=1000.0,
                               g_x=0.0, g_y=0.0, tau=1.0, omega=1.7, epsilon=1e-4, max_it=max_it, n_print=1)


# ----------------------------------------------------------------------------- clocks
class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thr = threading.Thread(target=self._pump, daemon=True)
            self.thr.start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), line.strip()))

    def stop(self, t0, t1):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, smax, reasons, pw = [], None, set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ts, line in self.rows:
            if ts < t0 - 0.05 or ts > t1 + 0.05:
                continue
            f = [x.strip() for x in line.split(",")]
            try:
                sm.append(float(f[0])); smax = float(f[1]); pw.append(float(f[2]))
            except Exception:
                continue
            for n, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": smax,
                "power_w_max": max(pw) if pw else None, "samples": len(sm), "reasons": sorted(reasons)}


def host_cpu():
    model = "unknown"
    try:
        for line in open("/proc/cpuinfo"):
            if line.lower().startswith("model name"):
                model = line.split(":", 1)[1].strip()
                break
    except Exception:
        pass
    return {"cpu_model": model, "nproc": os.cpu_count()}


def bind_to_gpu_numa(local):
    """One process per GPU: run this process on the CPUs of the GPU's NUMA node, so that the pinned host buffers it
    allocates afterwards are first-touched there (without a binding every rank's 3.2 GB per step
    of pinned traffic may cross the socket interconnect).  Returns what was done, for the JSON line."""
    try:
        import torch
        pr = torch.cuda.get_device_properties(local)
        bus = "%04x:%02x:%02x.0" % (getattr(pr, "pci_domain_id", 0), pr.pci_bus_id, pr.pci_device_id)
        node = int(open(f"/sys/bus/pci/devices/{bus}/numa_node").read())
        if node < 0:
            return {"numa_node": None, "note": "the platform reports no NUMA affinity for the GPU"}
        cpus = set()
        for part in open(f"/sys/devices/system/node/node{node}/cpulist").read().strip().split(","):
            a, _, b = part.partition("-")
            cpus.update(range(int(a), int(b or a) + 1))
        allowed = cpus & os.sched_getaffinity(0)
        if not allowed:
            return {"numa_node": node, "note": "no allowed CPU on that node"}
        os.sched_setaffinity(0, allowed)
        return {"numa_node": node, "cpus": len(allowed), "pci": bus}
    except Exception as e:               # binding is an optimisation, never a requirement
        return {"numa_node": None, "note": f"not bound: {type(e).__name__}: {e}"}


# ----------------------------------------------------------------------------- CPU reference leg
def cpu_reference_sor(size, iters_per_step, steps, warmup):
    """Times the reference's own serial SOR (src/serial/integration.c:129-173, compiled unmodified
    into oracle/_ref/libnsref.so; falls back to the oracle port if that build is absent) on the
    bench workload: p = 0, RHS = the cavity's first-step RHS, `iters_per_step` sweeps per step."""
    from oracle import binding as ob
    P = workload_params(ob, size, iters_per_step)
    L = ob.lib()
    ni = nj = size
    dx, dy = P.a / ni, P.b / nj
    u, v, F, G, RHS, p, res = (ob.zeros(P) for _ in range(7))
    dt, gamma = C.c_double(), C.c_double()
    L.orc_dt_gamma(C.byref(P), dx, dy, 0.0, 0.0, C.byref(dt), C.byref(gamma))
    L.orc_apply_bc(C.byref(P), 0.0, ob._ptr(u), ob._ptr(v))
    L.orc_FG(ob._ptr(F), ob._ptr(G), ob._ptr(u), ob._ptr(v), ni, nj, P.Re, 0.0, 0.0, dt.value, dx, dy, gamma.value)
    L.orc_RHS(ob._ptr(RHS), ob._ptr(F), ob._ptr(G), ni, nj, dt.value, dx, dy)
    kind = "reference" if ob.have_ref() else "port"
    if kind == "reference":
        R = ob.ref()
        pr, rr, hr = ob.rows(p), ob.rows(res), ob.rows(RHS)
        run = lambda: R.SOR(pr, ni, nj, dx, dy, rr, hr, P.omega, 0.0, iters_per_step)   # eps=0: no early exit
    else:
        run = lambda: ob.sor_lex(p, RHS, dx, dy, P.omega, 0.0, iters_per_step)
    for _ in range(warmup):
        run()
    t0 = time.perf_counter()
    for _ in range(steps):
        run()
    el = time.perf_counter() - t0
    lups = float(steps) * iters_per_step * ni * nj
    return {"value": lups / el / 1e6, "unit": "MLUPS", "cores": 1, "kind": kind, "host": host_cpu(),
            "sample": f"{steps} x SOR({iters_per_step} lexicographic sweeps, residual every sweep) on the "
                      f"{size}x{size} cavity RHS, {'oracle/_ref/libnsref.so (unmodified reference serial SOR)' if kind == 'reference' else 'oracle port'}, "
                      f"single thread (the reference is single-threaded), {el:.1f} s",
            "seconds": el, "ms_per_step": el / steps * 1e3}



# ----------------------------------------------------------------------------- parity gate
def parity_gate(m, rank, world, local, dist, pass_k, chunk_rows, ni=2048, rows_per_gpu=2048, iters=30):
    """Runs BEFORE the timed region, at every N, through the SAME engine the timing uses (persistent sweep, the
    multi-slab kernel with in-flight halo flags at N > 1, the timed run's iterations per pass and chunk height): one whole time step of a
    ni x (rows_per_gpu * N) cavity with max_it = `iters` (BC, F/G, RHS, `iters` red+black+residual iterations,
    velocity update) and `iters` more bare iterations.  Rank 0 then solves the same GLOBAL problem on one slab and
    with the CPU oracle (all host cores; the oracle is the checker, never the thing timed) and compares, bit for bit,
    sha256 of every rank's rows of u, v, p, plus dt and the two residuals (hex).  A mismatch ends the run non-zero."""
    import hashlib
    import torch
    nj = rows_per_gpu * world
    P = m.make_params(problem=1, f=1.0, i_max=ni, j_max=nj, a=1.0, b=float(nj) / float(ni), T=1.0e9, Re=1000.0,
                      g_x=0.0, g_y=0.0, tau=1.0, omega=1.7, epsilon=1e-4, max_it=iters, n_print=1)
    ch = chunk_rows

    def digest(rows):           # rows: C-contiguous (n_rows, ni + 2) block, x fastest
        return hashlib.sha256(np.ascontiguousarray(rows).tobytes()).hexdigest()

    def run(s, j_lo, nj_loc):
        info = s.step(0.0)
        res2 = s.sor_iterations(iters)
        out = {"dt": float(info["dt"]).hex(), "res": float(info["res"]).hex(), "res2": float(res2).hex(),
               "iters": info["sor_iters"], "j_lo": j_lo, "nj": nj_loc}
        for name, f in (("u", m.U), ("v", m.V), ("p", m.P)):
            a = s.download(f, layout=m.LAYOUT_XFAST)
            out[name] = digest(a[1:nj_loc + 1])
        return out

    t0 = time.time()
    if world == 1:
        mine = None
    else:
        s = m.SlabSolver(P, rank, world, m.torch_all_gather(), device=local, chunk_rows=ch, pass_k=pass_k)
        mine = run(s, s.j_lo, s.nj)
        engine = "sor_persist_kernel<*, MULTI=1, K=%d>, %d slabs, chunk_rows %d" % (s.sor_plan()["pass_k"], world, s.chunk_rows)
        dist.barrier()
        s.close()
        blobs = m.torch_all_gather()(json.dumps(mine).encode().ljust(1024))
        parts = [json.loads(b.decode().strip()) for b in blobs]
    result = None
    if rank == 0:
        # the same global problem on ONE slab (single-slab instantiation of the same sweep)
        with m.Solver(P, device=local, sor_engine=m.SOR_STREAM, chunk_rows=ch, pass_k=pass_k) as s1:
            k1 = s1.sor_plan()["pass_k"]
            info = s1.step(0.0)
            res2 = s1.sor_iterations(iters)
            one = {"dt": float(info["dt"]).hex(), "res": float(info["res"]).hex(), "res2": float(res2).hex(),
                   "iters": info["sor_iters"]}
            F1 = {name: s1.download(f, layout=m.LAYOUT_XFAST) for name, f in (("u", m.U), ("v", m.V), ("p", m.P))}
        if world == 1:
            engine = "sor_persist_kernel<*, MULTI=0, K=%d>, chunk_rows %d" % (k1, ch)
            parts = []
        ok_slabs = True
        for q in parts:
            for k in ("dt", "res", "res2", "iters"):
                ok_slabs = ok_slabs and (q[k] == one[k])
            for name in ("u", "v", "p"):
                ok_slabs = ok_slabs and (q[name] == digest(F1[name][q["j_lo"] + 1:q["j_lo"] + q["nj"] + 1]))
        # CPU oracle: the same step + iterations, all host cores for the sweeps
        from oracle import binding as ob
        L = ob.lib()
        oP = ob.make_params(**{k: getattr(P, k) for k, _ in ob.Params._fields_})
        dx, dy = P.a / ni, P.b / nj
        u, v, F, G, R, pp = (ob.zeros(oP) for _ in range(6))
        dt, gamma = C.c_double(), C.c_double()
        L.orc_dt_gamma(C.byref(oP), dx, dy, 0.0, 0.0, C.byref(dt), C.byref(gamma))
        L.orc_apply_bc(C.byref(oP), 0.0, ob._ptr(u), ob._ptr(v))
        L.orc_FG(ob._ptr(F), ob._ptr(G), ob._ptr(u), ob._ptr(v), ni, nj, P.Re, 0.0, 0.0, dt.value, dx, dy, gamma.value)
        L.orc_RHS(ob._ptr(R), ob._ptr(F), ob._ptr(G), ni, nj, dt.value, dx, dy)
        _, _, ores = ob.sor_rb_omp(pp, R, dx, dy, P.omega, 0.0, iters)
        L.orc_update_uv(ob._ptr(u), ob._ptr(v), ob._ptr(F), ob._ptr(G), ob._ptr(pp), ni, nj, dt.value, dx, dy)
        _, _, ores2 = ob.sor_rb_omp(pp, R, dx, dy, P.omega, 0.0, iters)
        ok_orc = (float(dt.value).hex() == one["dt"]) and one["iters"] == iters
        ok_orc = ok_orc and abs(float.fromhex(one["res"]) - ores) <= 1e-12 * ores and abs(float.fromhex(one["res2"]) - ores2) <= 1e-12 * ores2
        for name, arr in (("u", u), ("v", v), ("p", pp)):
            ok_orc = ok_orc and np.array_equal(F1[name][1:nj + 1, 1:ni + 1], arr.T[1:nj + 1, 1:ni + 1])
        result = {"checked": True, "ok": bool(ok_slabs and ok_orc), "engine": engine,
                  "case": "%dx%d cavity, 1 whole step (max_it %d) + %d iterations" % (ni, nj, iters, iters),
                  "vs": (["single-slab run of the same global problem (sha256 of u, v, p rows per rank, dt / residual hex)"] if world > 1 else []) +
                        ["oracle/liboracle.so (orc_FG, orc_RHS, orc_SOR_rb_omp, orc_update_uv): u, v, p bit-exact, residuals 1e-12"],
                  "ok_vs_single_slab": bool(ok_slabs) if world > 1 else None, "ok_vs_oracle": bool(ok_orc),
                  "seconds": round(time.time() - t0, 1)}
    if world > 1:
        flag = torch.tensor([1 if (result is None or result["ok"]) else 0], device="cuda")
        dist.broadcast(flag, 0)
        if int(flag.item()) == 0 and rank != 0:
            raise SystemExit(3)
    return result

# ----------------------------------------------------------------------------- main
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--size", type=int, default=8192, help="i_max (and j_max per GPU with weak scaling)")
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"],
                    help="N > 1: weak = size x (size*N) cells, one size^2 slab per GPU; strong = size^2 cut into N slabs")
    ap.add_argument("--max-it", type=int, default=500)
    ap.add_argument("--e2e-steps", type=int, default=2)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--engine", type=int, default=0)
    ap.add_argument("--chunk-rows", type=int, default=0)
    ap.add_argument("--pass-k", type=int, default=0, help="SOR iterations per pass over HBM: 1, 2, or 0 = planner's choice")
    ap.add_argument("--no-parity", action="store_true", help="skip the parity gate that runs before the timed region")
    args = ap.parse_args()

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    size, max_it = args.size, args.max_it
    ni = size
    nj = size * world if (world > 1 and args.scaling == "weak") else size     # global rows
    cells = float(ni) * float(nj)                                             # global cells
    if world == 1:
        which = {8192: "configs[2]", 16384: "configs[3] grid on one GPU", 32768: "configs[4] box on one GPU"}.get(size, "other size")
        workload = (f"synthetic lid-driven cavity {size}x{size} FP64, Re=1000, omega=1.7, eps=1e-4, "
                    f"max_it={max_it} (BASELINE.json {which})")
        par = "single GPU"
    elif args.scaling == "weak":
        workload = (f"synthetic lid-driven cavity {ni}x{nj} FP64 ({size}^2 cells per GPU, box {1}x{world}), Re=1000, "
                    f"omega=1.7, eps=1e-4, max_it={max_it} (BASELINE.json configs[4], weak scaling of configs[2])")
        par = f"y-slabs x{world}, halo rows pushed over NVLink inside the SOR kernel, one process per GPU"
    else:
        workload = (f"synthetic lid-driven cavity {ni}x{nj} FP64 cut into {world} y-slabs, Re=1000, omega=1.7, "
                    f"eps=1e-4, max_it={max_it} (BASELINE.json configs[3], strong scaling)")
        par = f"y-slabs x{world}, halo rows pushed over NVLink inside the SOR kernel, one process per GPU"
    config = {"workload": workload, "grid": [ni, nj], "max_it": max_it, "parallelism": par,
              "l2": f"inputs larger than L2 ({ni * (nj // world) * 8 / 1e6:.0f} MB per grid per GPU vs 126 MB L2); no flush"}
    scaling = "weak" if (world == 1 or args.scaling == "weak") else "strong"

    if args.impl == "reference":
        if rank != 0:
            return 0
        # bounded sample: 1 lexicographic sweep over the 8192^2 grid ~ 1.1 s on one core
        ips = 2 if size >= 4096 else max(2, int(2e8 / (float(size) * size)))
        b = cpu_reference_sor(size, ips, max(1, args.steps), max(1, min(args.warmup, 1)))
        line = {"impl": "reference", "metric": "sor_mlups", "value": b["value"], "unit": "MLUPS", "n_gpus": args.gpus,
                "steps": args.steps, "warmup": args.warmup, "ms_per_step": b["ms_per_step"], "higher_is_better": True,
                "scaling": scaling, "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": config,
                "cpu_baseline": {k: b[k] for k in ("value", "unit", "cores", "kind", "sample", "host")},
                "e2e": {"value": b["value"], "unit": "MLUPS", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                "gpu_launches": 0,
                "reference_arm_note": (f"the reference has no multi-GPU / multi-threaded path: this arm always times its serial "
                                       f"SOR on ONE host core on a {size}x{size} grid (bounded sample), whatever --gpus says; "
                                       f"at N > 1 the GPU arm's grid is {ni}x{nj}")}
        print(json.dumps(line))
        return 0

    import torch
    import torch.distributed as dist
    m = pkg()
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product has no CPU path")
    torch.cuda.set_device(local)
    binding = bind_to_gpu_numa(local) if world > 1 else None
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if world == 1:
            return x
        tt = torch.tensor([x], device="cuda", dtype=torch.float64)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        return float(tt.item())

    def sum_over_ranks(x):
        if world == 1:
            return x
        tt = torch.tensor([x], device="cuda", dtype=torch.float64)
        dist.all_reduce(tt, op=dist.ReduceOp.SUM)
        return float(tt.item())

    P = m.make_params(problem=1, f=1.0, i_max=ni, j_max=nj, a=1.0, b=float(nj) / float(ni), T=1.0e9, Re=1000.0,
                      g_x=0.0, g_y=0.0, tau=1.0, omega=1.7, epsilon=1e-4, max_it=max_it, n_print=1)
    if world == 1:
        s = m.Solver(P, device=local, sor_engine=args.engine, chunk_rows=args.chunk_rows, pass_k=args.pass_k)
        nj_local = nj
    else:
        # one slab per process; NCCL only carries the 256-byte slab descriptors, once.
        # band height: planned by the library for the slab height unless given (8192-row slabs: 124 rows)
        ch = args.chunk_rows
        s = m.SlabSolver(P, rank, world, m.torch_all_gather(), device=local, chunk_rows=ch, pass_k=args.pass_k)
        nj_local = s.nj
    plan = s.sor_plan()

    parity = {"checked": False}
    if not args.no_parity:
        parity = parity_gate(m, rank, world, local, dist, plan["pass_k"], plan["chunk_rows"])
        if rank == 0 and not parity["ok"]:
            print(json.dumps({"metric": "sor_mlups", "value": None, "n_gpus": world, "parity": parity,
                              "error": "parity gate failed: nothing was timed"}))
            sys.exit(3)
        barrier()
    cells_local = float(ni) * float(nj_local)
    stream = torch.cuda.ExternalStream(s.stream, device=torch.device("cuda", local))

    # ---- device-resident steps
    t = 0.0
    for _ in range(args.warmup):
        info = s.step(t); t += info["dt"]
    sampler = ClockSampler(local)
    sampler.start()
    time.sleep(0.3)
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    w0 = time.time()
    l0 = s.launch_count()
    e0.record(stream)
    iters = 0
    for _ in range(args.steps):
        info = s.step(t); t += info["dt"]; iters += info["sor_iters"]
    e1.record(stream)
    barrier()
    w1 = time.time()
    launches = int(sum_over_ranks(float(s.launch_count() - l0)))
    ms = max_over_ranks(e0.elapsed_time(e1))
    clocks = sampler.stop(w0, w1)
    total_lups = float(iters) * cells                      # every iteration updates every cell of the global grid once
    value = total_lups / (ms * 1e-3) / 1e6

    # ---- dominant kernel alone: average launch duration over a fixed batch of iterations
    barrier()
    n_k = 100
    s.sor_iterations(10)
    barrier()
    k0, k1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    k0.record(stream)
    s.sor_iterations(n_k)
    k1.record(stream)
    torch.cuda.synchronize()
    k_ms = max_over_ranks(k0.elapsed_time(k1) / n_k)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    achieved = 24.0 * cells_local / (k_ms * 1e-3) / 1e9
    traffic, traffic_src = None, None
    try:
        tj = json.load(open(os.path.join(ROOT, "profiles", "sor_stream_traffic.json")))
        traffic = tj["dram_bytes_per_iteration"][str(plan["pass_k"])] * n_k
        traffic_src = "static, not measured in this run: " + tj["source"][str(plan["pass_k"])]
    except Exception:
        pass
    roofline = {"bound": "hbm", "kernel": "sor_persist_kernel<K=%d>" % plan["pass_k"], "achieved": achieved, "peak": peak, "unit": "GB/s",
                "frac": achieved / peak, "traffic": traffic if (world == 1 and size == 8192) else None,
                "traffic_source": traffic_src if (world == 1 and size == 8192) else None,
                "peak_source": "MEASURED_PEAKS.json hbm_gbs (of measured)" if peaks else "6650 GB/s (of fallback)",
                "iterations_per_launch": n_k, "iterations_per_pass": plan["pass_k"], "chunk_rows": plan["chunk_rows"],
                "algorithmic_bytes_per_launch": 24.0 * cells_local * n_k,
                "launch_ms": k_ms * n_k, "ms_per_iteration": k_ms,
                "kernel_mlups": cells_local / (k_ms * 1e-3) / 1e6,
                "note": "per GPU: one launch = %d red+black+residual iterations over this GPU's cells (persistent sweep, %d "
                        "iteration(s) per pass over HBM); `achieved` counts the ALGORITHMIC 24 B per cell per iteration (p in, RHS in, "
                        "p out), so with two iterations per pass (temporal blocking) it can exceed the HBM peak -- `traffic` = "
                        "DRAM bytes of that launch from ncu" % (n_k, plan["pass_k"])}

    # ---- end to end: host grids in, host grids out, pinned host memory, copies inside the timed region
    e2e = None
    if not args.no_e2e:
        L = m.lib()
        tt_, it2 = 0.0, 0
        if world == 1:
            # the reference-shaped host-grid call: nsb_timestep(double** u, v, p)
            hu = torch.zeros((ni + 1, nj + 2), dtype=torch.float64).pin_memory()
            hv = torch.zeros((ni + 2, nj + 1), dtype=torch.float64).pin_memory()
            hp = torch.zeros((ni + 2, nj + 2), dtype=torch.float64).pin_memory()
            ur, vr, pr = m.rows(hu.numpy()), m.rows(hv.numpy()), m.rows(hp.numpy())
            null = C.cast(None, m.dbl_pp)
            info = m.StepInfo()

            def one_step(tq):
                rc = L.nsb_timestep(C.byref(P), tq, ur, vr, pr, null, null, null, C.byref(info))
                assert rc == 0, rc
                return info.dt, info.sor_iters
            bi = (hu.numel() + hv.numel() + hp.numel()) * 8
            api = "nsb_timestep(double** u, v, p) -- upload, whole step in HBM, download"
        else:
            # each rank keeps its slab (rows 0 .. nj_local+1) of u, v, p in pinned host memory
            hb = [torch.zeros((nj_local + 2, ni + 2), dtype=torch.float64).pin_memory() for _ in range(3)]
            ptrs = [C.cast(x.data_ptr(), m.dbl_p) for x in hb]
            fields = (m.U, m.V, m.P)

            def one_step(tq):
                for f, q in zip(fields, ptrs):
                    rc = L.nsb_upload(s._h, f, q, m.LAYOUT_XFAST)
                    assert rc == 0, rc
                inf = s.step(tq)
                for f, q in zip(fields, ptrs):
                    rc = L.nsb_download(s._h, f, q, m.LAYOUT_XFAST)
                    assert rc == 0, rc
                return inf["dt"], inf["sor_iters"]
            bi = sum(x.numel() for x in hb) * 8 * world
            api = "per rank: nsb_upload(u, v, p slab) -> nsb_step -> nsb_download(u, v, p slab)"
        dt_, _ = one_step(tt_)                                   # warm-up (allocates)
        tt_ += dt_
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.e2e_steps):
            dt_, k_ = one_step(tt_)
            tt_ += dt_; it2 += k_
        barrier()
        el = max_over_ranks(time.perf_counter() - t0)
        e2e = {"value": float(it2) * cells / el / 1e6, "unit": "MLUPS", "h2d_bytes_per_step": bi,
               "d2h_bytes_per_step": bi, "steps": args.e2e_steps, "ms_per_step": el / args.e2e_steps * 1e3, "api": api}
        if binding is not None:
            e2e["host_binding_rank0"] = binding
        if world == 1:
            L.nsb_host_api_reset()
    barrier()
    s.close()

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        b = cpu_reference_sor(size, 2 if size >= 4096 else max(2, int(2e8 / cells)), 8 if size >= 4096 else 5, 1)
        cpu = {k: b[k] for k in ("value", "unit", "cores", "kind", "sample", "host")}

    if rank == 0:
        line = {"metric": "sor_mlups", "value": value, "unit": "MLUPS", "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": scaling,
                "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": config,
                "timesteps_per_s": args.steps / (ms * 1e-3), "sor_iters_per_step": iters / max(1, args.steps),
                "hbm_gbs_whole_step": (80.0 + 24.0 * iters / max(1, args.steps)) * cells * args.steps / (ms * 1e-3) / 1e9,
                "clocks": clocks, "e2e": e2e, "gpu_launches": launches, "roofline": roofline, "cpu_baseline": cpu,
                "parity": parity}
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
