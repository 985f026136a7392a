#!/usr/bin/env python
"""Benchmark of the PPO_OBJ learner hot path (BASELINE.json metric).

    python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path
    python bench.py --impl reference --steps K --warmup W    # the reference's torch-CPU path

A "step" is one learner iteration over one synthetic rollout batch: bootstrap
value + GAE + the 4-epoch x 4-minibatch PPO update (cleanrl/ppo_atari_oc.py:532-617)
on Pong-shaped object vectors, 4096 envs x 128 steps per GPU (BASELINE config 2).
Prints ONE JSON line on rank 0.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
# stdout carries the ONE JSON line only: everything any library writes to fd 1 (NCCL prints its version banner there
# with printf) is sent to stderr, and the line itself goes to a private duplicate of the original stdout.
_JSON_FD = os.dup(1)
os.dup2(2, 1)


def emit(line):
    sys.stdout.flush()
    os.write(_JSON_FD, (json.dumps(line) + "\n").encode())


import numpy as np  # noqa: E402
import torch  # noqa: E402

ENC, DEC = (256, 512, 1024, 512), (512,)
FRAMES = 4
# dram__bytes_read.sum + dram__bytes_write.sum of one tc_gemm_kernel<FWD> launch (M=524288, N=512, K=1024) from the
# ncu --set full capture profiles/r01_tc8_ncu_full.txt (engine 8; 2.247 GB read + 1.044 GB written; algorithmic bytes of that
# launch: 3.22 GB = the A pair and the C pair once)
TRAFFIC_BYTES_PER_LAUNCH = 3291000000


def parse():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=3)
    p.add_argument("--warmup", type=int, default=3)
    p.add_argument("--impl", default="b200", choices=["b200", "reference"])
    p.add_argument("--game", default="pong", choices=["pong", "spaceinvaders", "breakout"])
    p.add_argument("--num-envs", type=int, default=4096, help="envs PER GPU (weak scaling)")
    p.add_argument("--num-steps", type=int, default=128)
    p.add_argument("--chunk-rows", type=int, default=32768)
    p.add_argument("--no-cpu-baseline", action="store_true")
    p.add_argument("--no-e2e", action="store_true")
    p.add_argument("--gae-sweep", default="4096,16384,65536")
    p.add_argument("--gemm", default="tc", choices=["fma", "tc", "tc-pair2", "tc-f32", "tc-bf16", "tc-3xtf32", "tc-fast"],
                   help="trunk GEMM engine: fp32 CUDA-core FMA | tcgen05, three FP16 passes on operands stored as fp16 (hi, lo) "
                        "pairs, one double-buffered accumulator for FWD / DGRAD (library default) | the same with split "
                        "accumulators everywhere | three FP16 passes on fp32-stored operands (in-kernel conversion) | TF32 main "
                        "product + two BF16 correction products | 3xTF32 | as tc-bf16 with one accumulator")
    return p.parse_args()


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        d = json.load(open(path))
        return dict(hbm=d["hbm_gbs"], bf16=d["bf16_tflops"], bf16_sustained=d.get("bf16_tflops_sustained", d["bf16_tflops"]),
                    source="measured (MEASURED_PEAKS.json)")
    return dict(hbm=6650.0, bf16=1590.0, bf16_sustained=1400.0, source="fallback (B200_PROFILING.md)")


def trunk_macs(D):
    m, w = 0, D
    for out in ENC:
        m += FRAMES * w * out
        w = out
    w *= FRAMES
    for out in DEC:
        m += w * out
        w = out
    return m


def gemm_flops_per_step(D, n_envs, n_steps, epochs=4):
    """Algorithmic trunk-GEMM FLOPs of one step: update fwd+dgrad+wgrad over
    `epochs` passes of the batch + the bootstrap get_value forward."""
    fwd = trunk_macs(D)
    upd = 3 * fwd - FRAMES * D * ENC[0]
    return 2.0 * upd * n_envs * n_steps * epochs + 2.0 * fwd * n_envs


def gae_bytes(T, N):
    return 20 * T * N + 8 * N


class ClockSampler:
    FIELDS = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.path = index, None, f"/tmp/ocppo_clocks_{os.getpid()}.csv"

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.FIELDS}", "--format=csv,noheader,nounits",
                                          "-lms", "200", "-i", str(self.index)], stdout=open(self.path, "w"),
                                         stderr=subprocess.DEVNULL)
        except OSError:
            self.proc = None

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, smax, power, reasons = [], [], [], set()
        names = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")
        for line in open(self.path):
            f = [x.strip() for x in line.split(",")]
            if len(f) < 8:
                continue
            try:
                sm.append(float(f[1])); smax.append(float(f[2])); power.append(float(f[3]))
            except ValueError:
                continue
            for name, val in zip(names, f[4:8]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        try:
            os.remove(self.path)
        except OSError:
            pass
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        return {"sm_mhz": statistics.median(sm), "sm_max_mhz": max(smax), "power_w_max": max(power),
                "samples": len(sm), "reasons": sorted(reasons)}


# --------------------------------------------------------------------------
# CPU arm: the reference's torch path (oracle restatement + OraclePPObj) on host cores
# --------------------------------------------------------------------------
def cpu_reference_step(D, A, n_envs, n_steps, sample_envs=128, mb_cap=4096, repeats=1):
    """Times GAE on the FULL (T, N) buffers and the update on a bounded sample
    (sample_envs x n_steps samples, one epoch of minibatches of <= mb_cap), then
    extrapolates the update linearly to 4 epochs of the full batch.
    Returns (samples_per_s, detail dict)."""
    from oracle import ref_ppo
    from oc_cleanrl_b200.synthetic import make_rollout
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    T = n_steps
    agent = ref_ppo.make_agent(FRAMES, D, A, ENC, DEC, seed=42)
    args = ref_ppo.PPOArgs(num_envs=n_envs, num_steps=T)
    g = torch.Generator().manual_seed(0)
    rewards, values = torch.randn(T, n_envs, generator=g), torch.randn(T, n_envs, generator=g)
    dones = (torch.rand(T, n_envs, generator=g) < 0.005).float()
    nv, nd = torch.randn(n_envs, generator=g), torch.zeros(n_envs)
    ref_ppo.ref_gae(rewards, values, dones, nv, nd, args.gamma, args.gae_lambda)          # warm-up
    t_gae = []
    for _ in range(5):
        t0 = time.perf_counter()
        ref_ppo.ref_gae(rewards, values, dones, nv, nd, args.gamma, args.gae_lambda)
        t_gae.append(time.perf_counter() - t0)
    t_gae = statistics.median(t_gae)
    data = make_rollout(T, sample_envs, FRAMES, D)
    n = T * sample_envs
    b_obs = data["obs"].reshape(n, FRAMES, D)
    with torch.no_grad():
        b_act, b_logp, _, b_val = agent.get_action_and_value(b_obs)
    b_adv, b_ret = torch.randn(n, generator=g), torch.randn(n, generator=g)
    batch = (b_obs, b_act, b_logp, b_adv, b_ret, b_val.reshape(-1))
    opt = ref_ppo.make_optimizer(agent, args)
    mb = min(mb_cap, n)
    ref_ppo.ref_minibatch_step(agent, opt, batch, np.arange(min(512, n)), args)           # warm-up
    t_upd = []
    for _ in range(repeats):
        perm = np.random.RandomState(0).permutation(n)
        t0 = time.perf_counter()
        for s in range(0, n, mb):
            ref_ppo.ref_minibatch_step(agent, opt, batch, perm[s:s + mb], args)
        t_upd.append(time.perf_counter() - t0)
    t_upd = statistics.median(t_upd)
    full_passes = args.update_epochs * T * n_envs
    t_step = t_gae + t_upd * (full_passes / n)
    sps = T * n_envs / t_step
    detail = dict(cores=cores, torch_threads=torch.get_num_threads(), gae_ms_full=t_gae * 1e3,
                  update_s_sample=t_upd, sample_passes=n, extrapolated=True,
                  sample=f"GAE full {T}x{n_envs}; update: 1 epoch over {sample_envs} envs x {T} steps "
                         f"({n} samples, minibatch {mb}) extrapolated linearly to {args.update_epochs} epochs x "
                         f"{T * n_envs} samples")
    return sps, detail


def run_reference(a):
    rank = int(os.getenv("RANK", "0"))
    if rank != 0:
        return
    from oc_cleanrl_b200.synthetic import GAME_SHAPES
    D, A = GAME_SHAPES[a.game]
    vals, detail = [], None
    for i in range(a.warmup + a.steps):
        sps, detail = cpu_reference_step(D, A, a.num_envs, a.num_steps)
        if i >= a.warmup:
            vals.append(sps)
    v = statistics.median(vals)
    line = {
        "impl": "reference", "metric": "PPO learner samples/sec (GAE+update)", "value": v, "unit": "samples/s",
        "n_gpus": a.gpus, "steps": a.steps, "warmup": a.warmup, "ms_per_step": 1e3 * a.num_envs * a.num_steps / v,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(a, 1),
        "cpu_baseline": {"value": v, "unit": "samples/s", "cores": detail["cores"], "kind": "port",
                         "sample": detail["sample"], "detail": detail},
        "e2e": {"value": v, "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)


def workload_config(a, world):
    return {"workload": f"synthetic {a.game}-shaped OCAtari obj rollouts, {a.num_envs} envs x {a.num_steps} steps per GPU, "
                        f"GAE + 4-epoch x 4-minibatch PPO update (BASELINE configs[1])",
            "num_envs_per_gpu": a.num_envs, "num_steps": a.num_steps, "update_epochs": 4, "num_minibatches": 4,
            "encoder_dims": list(ENC), "decoder_dims": list(DEC), "world_size": world, "gemm_engine": a.gemm,
            "parallelism": f"env-sharded dp{world}, one flat-gradient NCCL allreduce per minibatch",
            "l2": "working set (>=1.3 GB activations per chunk, 50 MB obs) exceeds the 126 MB L2; no flush needed; "
                  "the isolated GAE timing rotates through >L2 worth of buffers"}


# --------------------------------------------------------------------------
# GPU arm
# --------------------------------------------------------------------------
def time_gae_isolated(dev, T, N, iters=200, min_bytes=192 << 20):
    """Average GAE launch duration with CUDA events around a graph replay of `iters` launches, rotating through
    more than an L2's worth of distinct buffers so no launch finds its inputs in L2."""
    from oc_cleanrl_b200.learner import compute_gae
    per_set = gae_bytes(T, N)
    n_sets = max(2, int(np.ceil(min_bytes / per_set)))
    g = torch.Generator(device=dev).manual_seed(0)
    sets = []
    for _ in range(n_sets):
        sets.append(dict(r=torch.randn(T, N, device=dev, generator=g), v=torch.randn(T, N, device=dev, generator=g),
                         d=(torch.rand(T, N, device=dev, generator=g) < 0.005).float(),
                         nv=torch.randn(N, device=dev, generator=g), nd=torch.zeros(N, device=dev),
                         out=(torch.empty(T, N, device=dev), torch.empty(T, N, device=dev))))
    def launch(s):
        compute_gae(s["r"], s["v"], s["d"], s["nv"], s["nd"], 0.99, 0.95, out=s["out"])
    for i in range(max(3, n_sets)):
        launch(sets[i % n_sets])
    torch.cuda.synchronize(dev)
    # The launches are replayed from a CUDA graph: a Python+ctypes call costs more host time
    # than the kernel runs for at N <= 16384, and a host-bound loop would time the host.
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        for i in range(iters):
            launch(sets[i % n_sets])
    graph.replay()
    torch.cuda.synchronize(dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    graph.replay()
    e1.record()
    torch.cuda.synchronize(dev)
    us = e0.elapsed_time(e1) * 1e3 / iters
    del graph
    del sets
    return us


def time_policy_step(agent, dev, D, n_envs, iters=10):
    """Rollout inference (ppo_atari_oc.py:505-509): one fused trunk + head + Philox sample call over `n_envs`
    observations already on the device; average ms per call with CUDA events."""
    obs = torch.randint(0, 200, (n_envs, FRAMES, D), device=dev).float()
    torch.cuda.manual_seed(1)
    with torch.no_grad():
        for _ in range(3):
            agent.get_action_and_value(obs)
        torch.cuda.synchronize(dev)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            agent.get_action_and_value(obs)
        e1.record()
        torch.cuda.synchronize(dev)
    return e0.elapsed_time(e1) / iters


def run_b200(a):
    from oc_cleanrl_b200 import _abi
    from oc_cleanrl_b200 import distributed as ocd
    from oc_cleanrl_b200.architectures.ppo import PPObj
    from oc_cleanrl_b200.learner import LearnerArgs, PPOLearner, compute_gae, make_permutations
    from oc_cleanrl_b200.synthetic import GAME_SHAPES, make_rollout
    from oracle.ref_ppo import stub_envs          # tiny env stub only (3 attributes), no oracle compute

    world = ocd.world_from_env()
    if not torch.cuda.is_available():
        raise SystemExit("bench.py (impl b200) needs a CUDA device; there is no CPU fallback")
    dev = torch.device("cuda", world.local_rank)
    torch.cuda.set_device(dev)
    ocd.init_process_group(world, "nccl")
    lib = _abi.lib()
    assert lib.ocppo_set_gemm_mode({"fma": 0, "tc": 8, "tc-pair2": 6, "tc-f32": 5, "tc-bf16": 3, "tc-3xtf32": 1, "tc-fast": 4}[a.gemm]) == 0
    D, A = GAME_SHAPES[a.game]
    T, N = a.num_steps, a.num_envs
    pk = peaks()

    rank_seed = ocd.seed_rank(1, world.rank)
    agent = PPObj(stub_envs(FRAMES, D, A), dev, ENC, DEC).to(dev)
    ocd.reseed_sampling(rank_seed)
    torch.cuda.manual_seed(rank_seed)
    largs = LearnerArgs(num_envs=N, num_steps=T)
    learner = PPOLearner(agent, largs, world_size=world.world_size, chunk_rows=a.chunk_rows)

    # ---- synthetic rollout for this rank, staged in pinned host memory -----------------
    data = make_rollout(T, N, FRAMES, D, seed=1234, rank=world.rank)
    host = {k: v.pin_memory() for k, v in data.items()}
    dbuf = {k: v.to(dev, non_blocking=True) for k, v in host.items()}
    with torch.no_grad():       # what the rollout phase would have stored (ppo_atari_oc.py:505-509)
        action, logp, _, value = agent.get_action_and_value(dbuf["obs"].reshape(T * N, FRAMES, D))
    dbuf["actions"], dbuf["logprobs"], dbuf["values"] = action.reshape(T, N), logp.reshape(T, N), value.reshape(T, N)
    for k in ("actions", "logprobs", "values"):
        host[k] = dbuf[k].cpu().pin_memory()
    torch.cuda.synchronize(dev)

    def step_device(perms):
        adv, ret = learner.compute_gae(dbuf["rewards"], dbuf["values"], dbuf["dones"], dbuf["next_obs"], dbuf["next_done"])
        return learner.update(dbuf["obs"], dbuf["actions"], dbuf["logprobs"], adv, ret, dbuf["values"], permutations=perms)

    def barrier():
        if world.distributed:
            torch.distributed.barrier()
        torch.cuda.synchronize(dev)

    def max_over_ranks(ms):
        if not world.distributed:
            return ms
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
        return float(t.item())

    total = a.warmup + a.steps
    perms = [make_permutations(T * N, largs.update_epochs) for _ in range(total)]   # resident inputs of `value`
    for i in range(a.warmup):
        step_device(perms[i])
    barrier()
    sampler = ClockSampler(world.local_rank)
    if world.rank == 0:
        sampler.start()
    lib.ocppo_profile_enable(1)
    launches0 = lib.ocppo_launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    last = None
    for i in range(a.steps):
        last = step_device(perms[a.warmup + i])
    e1.record()
    barrier()
    ms_total = max_over_ranks(e0.elapsed_time(e1))
    launches = lib.ocppo_launch_count() - launches0
    prof_ms, prof_n = _abi.profile_read()
    lib.ocppo_profile_enable(0)
    clocks = sampler.stop() if world.rank == 0 else None
    ms_per_step = ms_total / a.steps
    value_sps = world.world_size * T * N / (ms_per_step * 1e-3)

    # ---- e2e: host buffers in, stats out, through the public API ------------------------
    e2e = None
    if not a.no_e2e:
        h2d_keys = ("obs", "actions", "logprobs", "rewards", "dones", "values", "next_obs", "next_done")
        def step_e2e():
            d = {k: host[k].to(dev, non_blocking=True) for k in h2d_keys}
            adv, ret = learner.compute_gae(d["rewards"], d["values"], d["dones"], d["next_obs"], d["next_done"])
            # update() shuffles on the host as the reference does (:558-561), copies each epoch's order H2D, and
            # reads the stats back (D2H) before returning
            return learner.update(d["obs"], d["actions"], d["logprobs"], adv, ret, d["values"])
        step_e2e()
        barrier()
        s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s0.record()
        for _ in range(a.steps):
            step_e2e()
        s1.record()
        barrier()
        ms_e2e = max_over_ranks(s0.elapsed_time(s1)) / a.steps
        h2d = sum(host[k].numel() * host[k].element_size() for k in h2d_keys) + largs.update_epochs * T * N * 8
        n_mb = largs.update_epochs * largs.num_minibatches
        e2e = {"value": world.world_size * T * N / (ms_e2e * 1e-3), "unit": "samples/s", "ms_per_step": ms_e2e,
               "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(n_mb * (_abi.OCPPO_NUM_STATS + 1) * 4)}

    if world.rank != 0:
        if world.distributed:
            torch.distributed.destroy_process_group()
        return

    # ---- rooflines (rank 0) ---------------------------------------------------------
    gemm_ms = (prof_ms["sgemm_fwd"] + prof_ms["sgemm_dgrad"] + prof_ms["sgemm_wgrad"]) / a.steps
    gemm_flops = gemm_flops_per_step(D, N, T, largs.update_epochs)
    gemm_tflops = gemm_flops / (gemm_ms * 1e-3) / 1e12 if gemm_ms > 0 else 0.0
    fma = C_double()
    lib.ocppo_measure_fma_peak(fma, None)
    tc = a.gemm != "fma"
    roofline = {"bound": "tensor", "achieved": gemm_tflops, "peak": pk["bf16_sustained"], "unit": "TFLOP/s",
                "frac": gemm_tflops / pk["bf16_sustained"], "traffic": TRAFFIC_BYTES_PER_LAUNCH if tc else None,
                "kernel": ("tc_gemm_kernel<FWD|DGRAD|WGRAD> (tcgen05.mma cta_group::2: " +
                           {"tc": "three FP16 products (hi.hi + hi.lo + lo.hi) on operands stored in HBM as power-of-two scaled fp16 (hi, lo) pairs, conversion-free six-stage pipeline",
                            "tc-pair2": "three FP16 products on fp16-pair stored operands, split accumulators",
                            "tc-f32": "three FP16 products on fp32-stored operands converted in the kernel",
                            "tc-bf16": "TF32 main product + two BF16 correction products",
                            "tc-fast": "TF32 main product + two BF16 correction products, one accumulator",
                            "tc-3xtf32": "TF32 main product + two TF32 correction products"}.get(a.gemm, "") +
                           ", TMA-fed, TMEM accumulators, bulk-tensor-store epilogue) + the skinny first-layer/head kernels"
                           if tc else "sgemm_kernel (fp32 CUDA-core FMA; fwd+dgrad+wgrad launches of the trunk)"),
                "peak_source": pk["source"] + ", bf16 dense sustained",
                "share_of_step": gemm_ms / ms_per_step,
                "launches_per_step": (prof_n["sgemm_fwd"] + prof_n["sgemm_dgrad"] + prof_n["sgemm_wgrad"]) / a.steps,
                "flops_per_step": gemm_flops,
                "achieved_is": "ALGORITHMIC fp32-equivalent FLOPs (2*M*N*K per GEMM, SURVEY 8d) / CUDA-event time of the GEMM launches"}
    if tc:
        # bf16-equivalent MMA work issued per algorithmic MAC: a TF32 pass costs 2 (half the bf16 rate), a BF16 pass 1
        units = {"tc": 3.0, "tc-pair2": 3.0, "tc-f32": 3.0, "tc-bf16": 4.0, "tc-fast": 4.0}.get(a.gemm, 6.0)
        roofline["tensor_pipe"] = {"bf16_equivalent_tflops_issued": units * gemm_tflops, "peak": pk["bf16_sustained"],
                                   "frac_of_peak": units * gemm_tflops / pk["bf16_sustained"],
                                   "note": (("fp32-grade accuracy costs three FP16 passes (hi.hi, hi.lo, lo.hi)" if a.gemm in ("tc", "tc-pair2", "tc-f32") else
                                             "fp32-grade accuracy costs one TF32 pass (2 bf16-equivalents) plus " +
                                             ("two BF16 correction passes" if a.gemm in ("tc-bf16", "tc-fast") else "two TF32 correction passes")) +
                                            f": {units:.0f} bf16-equivalent MACs per algorithmic MAC, so frac = achieved/peak is "
                                            f"bounded by 1/{units:.0f} = {1.0 / units:.3f} at full tensor-pipe occupancy")}
    else:
        roofline["note"] = "fp32-exact FMA path: the tensor roofline is the ceiling the tcgen05 (3xTF32) trunk is held to"
    roofline_fma = {"bound": "fma", "achieved": gemm_tflops, "peak": fma.value, "unit": "TFLOP/s",
                    "frac": gemm_tflops / fma.value if fma.value else None,
                    "peak_source": "ocppo_measure_fma_peak micro-benchmark in this run"}
    sweep = []
    for n in [int(x) for x in a.gae_sweep.split(",") if x]:
        us = time_gae_isolated(dev, T, n)
        gbs = gae_bytes(T, n) / (us * 1e-6) / 1e9
        sweep.append({"num_envs": n, "num_steps": T, "us_per_launch": us, "bytes": gae_bytes(T, n), "GB/s": gbs,
                      "frac": gbs / pk["hbm"]})
    own = next((s for s in sweep if s["num_envs"] == N), sweep[0] if sweep else None)
    roofline_gae = None
    if own:
        roofline_gae = {"bound": "hbm", "achieved": own["GB/s"], "peak": pk["hbm"], "unit": "GB/s", "frac": own["frac"],
                        "traffic": None, "kernel": "gae_scan_kernel", "peak_source": pk["source"], "sweep": sweep}
    kernel_ms = {k: v / a.steps for k, v in prof_ms.items() if v > 0}
    # rollout inference over the same sweep (BASELINE configs[4]); the trunk is 11.55 MFLOP per observation against
    # 116 bytes in and out, so it is held to the tensor roofline like the update, not to HBM
    inference = []
    units = {"tc": 3.0, "tc-pair2": 3.0, "tc-f32": 3.0, "tc-bf16": 4.0, "tc-fast": 4.0}.get(a.gemm, 6.0) if tc else None
    for n in [int(x) for x in a.gae_sweep.split(",") if x]:
        ms = time_policy_step(agent, dev, D, n)
        tf = 2.0 * trunk_macs(D) * n / (ms * 1e-3) / 1e12
        inference.append({"num_envs": n, "ms_per_call": ms, "obs_per_s": n / (ms * 1e-3), "TFLOP/s": tf,
                          "frac_of_bf16_peak": tf / pk["bf16_sustained"],
                          "tensor_pipe_frac": units * tf / pk["bf16_sustained"] if units else None})

    cpu_baseline = None
    if world.world_size == 1 and not a.no_cpu_baseline:
        sps, detail = cpu_reference_step(D, A, N, T)
        cpu_baseline = {"value": sps, "unit": "samples/s", "cores": detail["cores"], "kind": "port",
                        "sample": detail["sample"], "detail": detail}

    line = {
        "metric": "PPO learner samples/sec (GAE+update)", "value": value_sps, "unit": "samples/s",
        "n_gpus": world.world_size, "steps": a.steps, "warmup": a.warmup, "ms_per_step": ms_per_step,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(a, world.world_size),
        "roofline": roofline, "roofline_fma": roofline_fma, "roofline_gae": roofline_gae,
        "inference": inference, "kernel_ms_per_step": kernel_ms, "cpu_baseline": cpu_baseline, "e2e": e2e,
        "gpu_launches": int(launches), "clocks": clocks,
        "last_stats": {k: last[k] for k in ("loss", "pg_loss", "v_loss", "entropy", "approx_kl", "grad_norm")},
        "impl": "b200",
    }
    emit(line)
    if world.distributed:
        torch.distributed.destroy_process_group()


def C_double():
    import ctypes
    return ctypes.c_double(0.0)


if __name__ == "__main__":
    args = parse()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)
