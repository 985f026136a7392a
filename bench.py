#!/usr/bin/env python
"""Benchmark of the PPO_OBJ learner hot path (BASELINE.json metric).

    python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path
    python bench.py --impl reference --steps K --warmup W    # the reference's torch-CPU path

A "step" is one learner iteration over one synthetic rollout batch: bootstrap value + GAE + the 4-epoch x 4-minibatch
PPO update (cleanrl/ppo_atari_oc.py:532-617) on Pong-shaped object vectors, 4096 envs x 128 steps per GPU
(BASELINE configs[1]).  Prints ONE JSON line on rank 0.

What is timed how (b200 arm):
  * `value`      -- K steps on device-resident inputs, CUDA events, profiling OFF, max over ranks;
  * `e2e`        -- K steps through the public API from pinned host buffers (H2D of the rollout every step, host shuffle,
                    H2D of the permutations, D2H of the statistics);
  * `roofline`, `kernel_ms_per_step` -- a SEPARATE pass of --prof-steps steps with every launch bracketed by CUDA events
                    (plain launches, no graphs: a captured launch cannot carry events); the pass asserts that the
                    number of kernels timed equals the number launched before anything is reported;
  * `configs`    -- BASELINE configs[0] (8 envs x 128 steps) and configs[2] (SpaceInvaders / Breakout shapes), 3 steps each;
  * `strong`     -- (N > 1) the same K steps with 4096 envs in TOTAL, ppo_atari_multigpu.py:166-172's other reading.
The reference arm times GAE at full size plus --ref-minibatches TRUE-SIZE minibatches (131,072 samples) per step.
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
EPOCHS, MINIBATCHES = 4, 4
# ncu --set full captures of single launches (dram__bytes_read.sum + dram__bytes_write.sum), valid for the configuration
# they were taken on only; `roofline.traffic` is null for any other
NCU_TRAFFIC = {
    # tc_gemm_kernel<FWD>, trunk layer 3 (M=524288, N=512, K=1024) inside a full-size minibatch of the learner path, engine 8:
    # 2.153 GB read + 1.072 GB written; algorithmic 3.22 GB (the A pair and the C pair once)
    ("pong", 4096, 128, "tc"): {"bytes": 3225000000, "source": "profiles/r02_gemm_ncu_minibatch.txt (tc_gemm_kernel<FWD> layer 3, one launch)"},
}
# gae_tma_kernel, ncu --set full, cold L2 (profiles/r02_gae_ncu.txt): dram read + write per launch.  The outputs of the two
# smaller sizes are still in the 126 MB L2 when the kernel ends, so ncu sees (almost) no DRAM write for them.
NCU_TRAFFIC_GAE = {
    (4096, 128): {"bytes": 6337792, "source": "profiles/r02_gae_ncu.txt (reads 6.34 MB = the three inputs; the 4.19 MB of outputs stay in L2)"},
    (16384, 128): {"bytes": 25312512, "source": "profiles/r02_gae_ncu.txt (reads 25.31 MB; outputs stay in L2)"},
    (65536, 128): {"bytes": 124140032, "source": "profiles/r02_gae_ncu.txt (101.20 MB read + 22.94 MB written back before the kernel ended)"},
}
ENGINES = {"fma": 0, "tc": 8, "tc-pair2": 6, "tc-f32": 5, "tc-bf16": 3, "tc-3xtf32": 1, "tc-fast": 4}
ENGINE_TEXT = {
    "tc": "three FP16 products (hi.hi + hi.lo + lo.hi) on operands stored in HBM as power-of-two scaled fp16 (hi, lo) pairs, conversion-free six-stage pipeline",
    "tc-pair2": "three FP16 products on fp16-pair stored operands, split accumulators",
    "tc-f32": "three FP16 products on fp32-stored operands converted in the kernel",
    "tc-bf16": "TF32 main product + two BF16 correction products",
    "tc-fast": "TF32 main product + two BF16 correction products, one accumulator",
    "tc-3xtf32": "TF32 main product + two TF32 correction products"}
PASS_UNITS = {"tc": 3.0, "tc-pair2": 3.0, "tc-f32": 3.0, "tc-bf16": 4.0, "tc-fast": 4.0, "tc-3xtf32": 6.0}


def parse():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=3)
    p.add_argument("--warmup", type=int, default=3)
    p.add_argument("--impl", default="b200", choices=["b200", "reference"])
    p.add_argument("--game", default="pong", choices=["pong", "spaceinvaders", "breakout"])
    p.add_argument("--num-envs", type=int, default=4096, help="envs PER GPU (weak scaling); total envs with --scaling strong")
    p.add_argument("--num-steps", type=int, default=128)
    p.add_argument("--scaling", default="weak", choices=["weak", "strong"])
    p.add_argument("--encoder-dims", default=",".join(map(str, ENC)))
    p.add_argument("--decoder-dims", default=",".join(map(str, DEC)))
    p.add_argument("--chunk-rows", type=int, default=131072,
                   help="samples per forward/backward chunk of the update (workspace size); 131072 = one chunk per minibatch")
    p.add_argument("--prof-steps", type=int, default=3)
    p.add_argument("--ref-minibatches", type=int, default=2, help="true-size minibatches timed per reference step")
    p.add_argument("--no-cpu-baseline", action="store_true")
    p.add_argument("--no-e2e", action="store_true")
    p.add_argument("--no-configs", action="store_true")
    p.add_argument("--no-strong", action="store_true")
    p.add_argument("--no-graphs", action="store_true", help="plain launches instead of CUDA-graph replay")
    p.add_argument("--gae-sweep", default="4096,16384,65536")
    p.add_argument("--gemm", default="tc", choices=list(ENGINES),
                   help="trunk GEMM engine: fp32 CUDA-core FMA | tcgen05, three FP16 passes on operands stored as fp16 (hi, lo) "
                        "pairs, one double-buffered accumulator for FWD / DGRAD (library default) | the same with split "
                        "accumulators everywhere | three FP16 passes on fp32-stored operands (in-kernel conversion) | TF32 main "
                        "product + two BF16 correction products | 3xTF32 | as tc-bf16 with one accumulator")
    a = p.parse_args()
    a.enc = tuple(int(x) for x in a.encoder_dims.split(",") if x)
    a.dec = tuple(int(x) for x in a.decoder_dims.split(",") if x)
    return a


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        d = json.load(open(path))
        return dict(hbm=d["hbm_gbs"], bf16=d["bf16_tflops"], bf16_sustained=d.get("bf16_tflops_sustained", d["bf16_tflops"]),
                    source="measured (MEASURED_PEAKS.json)")
    return dict(hbm=6650.0, bf16=1590.0, bf16_sustained=1400.0, source="fallback (B200_PROFILING.md)")


def trunk_macs(D, enc=ENC, dec=DEC):
    m, w = 0, D
    for out in enc:
        m += FRAMES * w * out
        w = out
    w *= FRAMES
    for out in dec:
        m += w * out
        w = out
    return m


def gemm_flops_per_step(D, n_envs, n_steps, enc=ENC, dec=DEC, epochs=EPOCHS):
    """Algorithmic trunk-GEMM FLOPs of one step: update fwd+dgrad+wgrad over
    `epochs` passes of the batch + the bootstrap get_value forward."""
    fwd = trunk_macs(D, enc, dec)
    upd = 3 * fwd - FRAMES * D * enc[0]
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
# CPU arm: the reference's torch path (oracle restatement of the loop around the reference's model class) on host cores
# --------------------------------------------------------------------------
class CpuReference:
    """GAE on the FULL (T, N) buffers and the update on TRUE-SIZE minibatches (batch_size // 4 samples of the full
    T x N rollout, ppo_atari_oc.py:562-610), a bounded number of them per step.  One whole update is
    update_epochs x num_minibatches = 16 such minibatches; the throughput of a step scales the timed minibatches to 16."""

    def __init__(self, D, A, n_envs, n_steps, enc=ENC, dec=DEC):
        from oracle import ref_ppo
        from oc_cleanrl_b200.synthetic import make_rollout
        self.ref = ref_ppo
        self.cores = os.cpu_count() or 1
        torch.set_num_threads(self.cores)
        self.T, self.N, self.n = n_steps, n_envs, n_steps * n_envs
        self.agent = ref_ppo.make_agent(FRAMES, D, A, enc, dec, seed=42)
        self.args = ref_ppo.PPOArgs(num_envs=n_envs, num_steps=n_steps)
        self.opt = ref_ppo.make_optimizer(self.agent, self.args)
        g = torch.Generator().manual_seed(0)
        data = make_rollout(n_steps, n_envs, FRAMES, D)
        self.gae_in = (data["rewards"], torch.randn(n_steps, n_envs, generator=g), data["dones"],
                       torch.randn(n_envs, generator=g), data["next_done"])
        # stored policy outputs of the rollout phase: a fresh PPObj is a near-uniform policy (actor gain 0.01), so the old
        # log-probs sit at log(1/A) +- 0.05 (a full CPU rollout forward to produce them would cost 20 s of set-up)
        rs = np.random.RandomState(1)
        self.batch = (data["obs"].reshape(self.n, FRAMES, D), torch.as_tensor(rs.randint(0, A, self.n)),
                      torch.as_tensor(-np.log(A) + 0.05 * rs.randn(self.n), dtype=torch.float32),
                      torch.randn(self.n, generator=g), torch.randn(self.n, generator=g), self.gae_in[1].reshape(-1).clone())
        self.mb = self.args.minibatch_size
        self.perm = np.random.RandomState(0).permutation(self.n)
        self.cursor = 0

    def gae(self):
        t0 = time.perf_counter()
        self.ref.ref_gae(*self.gae_in, self.args.gamma, self.args.gae_lambda)
        return time.perf_counter() - t0

    def minibatches(self, count):
        t0 = time.perf_counter()
        for _ in range(count):
            if self.cursor + self.mb > self.n:
                self.cursor = 0
                np.random.RandomState(self.cursor + 1).shuffle(self.perm)
            inds = self.perm[self.cursor:self.cursor + self.mb]
            self.cursor += self.mb
            self.ref.ref_minibatch_step(self.agent, self.opt, self.batch, inds, self.args)
        return time.perf_counter() - t0

    def step(self, n_mb):
        """-> (wall seconds actually spent, samples/s of a whole step with the update scaled to 16 minibatches)"""
        t_gae = self.gae()
        t_upd = self.minibatches(n_mb)
        total_mb = self.args.update_epochs * self.args.num_minibatches
        t_step = t_gae + t_upd * (total_mb / n_mb)
        return t_gae + t_upd, self.n / t_step, t_gae, t_upd

    def describe(self, n_mb):
        total_mb = self.args.update_epochs * self.args.num_minibatches
        return (f"per step: GAE on the full {self.T}x{self.N} buffers + {n_mb} TRUE-SIZE minibatches ({self.mb} samples each, "
                f"forward + backward + clip + Adam on the {self.n}-sample rollout); samples/s = {self.n} / "
                f"(t_gae + t_minibatches x {total_mb}/{n_mb}), i.e. the update time scaled x{total_mb / n_mb:g} to the {total_mb} "
                f"minibatches of one iteration; ms_per_step is the wall time actually spent per step")


def run_reference(a):
    rank = int(os.getenv("RANK", "0"))
    if rank != 0:
        return
    from oc_cleanrl_b200.synthetic import GAME_SHAPES
    D, A = GAME_SHAPES[a.game]
    n_envs = a.num_envs
    cpu = CpuReference(D, A, n_envs, a.num_steps, a.enc, a.dec)
    n_mb = max(1, a.ref_minibatches)
    for _ in range(a.warmup):                 # untimed; one minibatch each is enough to warm the thread pool / allocator
        cpu.gae()
        cpu.minibatches(1)
    walls, vals, gaes, upds = [], [], [], []
    for _ in range(a.steps):
        wall, sps, t_gae, t_upd = cpu.step(n_mb)
        walls.append(wall); vals.append(sps); gaes.append(t_gae); upds.append(t_upd)
    v = statistics.median(vals)
    total_mb = EPOCHS * MINIBATCHES
    detail = dict(cores=cpu.cores, torch_threads=torch.get_num_threads(), gae_ms_full=1e3 * statistics.median(gaes),
                  s_per_true_size_minibatch=statistics.median(upds) / n_mb, minibatches_timed_per_step=n_mb,
                  minibatches_timed_total=n_mb * a.steps, extrapolation_factor=total_mb / n_mb,
                  full_step_s_estimate=cpu.n / v)
    line = {
        "impl": "reference", "metric": "PPO learner samples/sec (GAE+update)", "value": v, "unit": "samples/s",
        "n_gpus": a.gpus, "steps": a.steps, "warmup": a.warmup, "ms_per_step": 1e3 * statistics.mean(walls),
        "higher_is_better": True, "scaling": a.scaling, "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(a, 1, n_envs),
        "cpu_baseline": {"value": v, "unit": "samples/s", "cores": cpu.cores, "kind": "port",
                         "sample": cpu.describe(n_mb), "detail": detail},
        "e2e": {"value": v, "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)


def workload_config(a, world, n_envs, extra=None):
    cfg = {"workload": f"synthetic {a.game}-shaped OCAtari obj rollouts, {n_envs} envs x {a.num_steps} steps per GPU, "
                       f"GAE + {EPOCHS}-epoch x {MINIBATCHES}-minibatch PPO update (BASELINE configs[1])",
           "num_envs_per_gpu": n_envs, "num_steps": a.num_steps, "update_epochs": EPOCHS, "num_minibatches": MINIBATCHES,
           "encoder_dims": list(a.enc), "decoder_dims": list(a.dec), "world_size": world, "gemm_engine": a.gemm,
           "parallelism": f"env-sharded dp{world}, one flat-gradient NCCL allreduce per minibatch",
           "l2": "working set (>=1.3 GB activations per chunk, 50 MB obs) exceeds the 126 MB L2; no flush needed; "
                 "the isolated GAE timing rotates through >L2 worth of buffers"}
    if extra:
        cfg.update(extra)
    return cfg


# --------------------------------------------------------------------------
# GPU arm
# --------------------------------------------------------------------------
def graph_timed(dev, launch, iters):
    """Average duration of `launch(i)` with CUDA events around a graph replay of `iters` launches (a Python + ctypes call
    costs more host time than these kernels run for, so a host-driven loop would time the host)."""
    for i in range(3):
        launch(i)
    torch.cuda.synchronize(dev)
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        for i in range(iters):
            launch(i)
    graph.replay()
    torch.cuda.synchronize(dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    graph.replay()
    e1.record()
    torch.cuda.synchronize(dev)
    us = e0.elapsed_time(e1) * 1e3 / iters
    del graph
    return us


def time_gae_isolated(dev, T, N, iters=200, min_bytes=192 << 20):
    """-> (us per GAE launch, us per same-bytes device copy, us per GAE launch in a chain of programmatic dependent
    launches -- how the learner launches it behind the bootstrap value kernel): all rotate through more than an L2's worth of distinct
    buffers so no launch finds its inputs in L2.  The copy (5 x T x N floats read + written = the GAE's algorithmic bytes)
    is the floor of any kernel that moves these bytes once at this size: launch + one DRAM round trip."""
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

    def launch(i):
        s = sets[i % n_sets]
        compute_gae(s["r"], s["v"], s["d"], s["nv"], s["nd"], 0.99, 0.95, out=s["out"])
    us = graph_timed(dev, launch, iters)

    def launch_overlapped(i):          # each launch a programmatic dependent of the one before it (other buffers)
        s = sets[i % n_sets]
        compute_gae(s["r"], s["v"], s["d"], s["nv"], s["nd"], 0.99, 0.95, out=s["out"], overlap_previous=True)
    us_pdl = graph_timed(dev, launch_overlapped, iters)
    # same bytes through cudaMemcpyAsync: 2.5 x T x N floats copied (read + write = 20 T N bytes)
    half = (5 * T * N) // 2
    flat = [torch.cat([s["r"].reshape(-1), s["v"].reshape(-1), s["d"].reshape(-1)])[:half].contiguous() for s in sets]
    dst = [torch.empty_like(f) for f in flat]

    def copy(i):
        dst[i % n_sets].copy_(flat[i % n_sets], non_blocking=True)
    us_copy = graph_timed(dev, copy, iters)
    del sets, flat, dst
    return us, us_copy, us_pdl


def time_policy_step(agent, dev, D, n_envs, iters=20):
    """Rollout inference (ppo_atari_oc.py:505-509): one fused trunk + head + Philox sample call over `n_envs`
    observations already on the device.  -> (ms per call through get_action_and_value, ms per call through
    policy_step_into(graph=True), the rollout collector's route: persistent buffers, CUDA-graph replay)."""
    obs = torch.randint(0, 200, (n_envs, FRAMES, D), device=dev).float()
    torch.cuda.manual_seed(1)
    out = (torch.empty(n_envs, dtype=torch.int64, device=dev), torch.empty(n_envs, device=dev),
           torch.empty(n_envs, device=dev), torch.empty(n_envs, device=dev))
    res = []
    with torch.no_grad():
        for call in (lambda: agent.get_action_and_value(obs), lambda: agent.policy_step_into(obs, *out, graph=True)):
            for _ in range(3):
                call()
            torch.cuda.synchronize(dev)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(iters):
                call()
            e1.record()
            torch.cuda.synchronize(dev)
            res.append(e0.elapsed_time(e1) / iters)
    return res


class Workload:
    """One (game shape, envs per GPU, steps) learner set-up on this rank: agent, fused learner, a synthetic rollout in pinned
    host memory and in persistent device buffers (as the rollout collector keeps them), and the stored policy outputs."""

    def __init__(self, a, world, dev, game, n_envs, T, enc, dec, chunk_rows, use_graphs, seed_rank):
        from oc_cleanrl_b200 import distributed as ocd
        from oc_cleanrl_b200.architectures.ppo import PPObj
        from oc_cleanrl_b200.learner import LearnerArgs, PPOLearner
        from oc_cleanrl_b200.synthetic import GAME_SHAPES, make_rollout, stub_envs
        self.world, self.dev, self.T, self.N = world, dev, T, n_envs
        self.D, self.A = GAME_SHAPES[game]
        torch.manual_seed(1)                   # common seed: identical initial weights on every rank (ppo_atari_multigpu.py:211)
        self.agent = PPObj(stub_envs(FRAMES, self.D, self.A), dev, enc, dec).to(dev)
        ocd.reseed_sampling(seed_rank)
        torch.cuda.manual_seed(seed_rank)
        self.largs = LearnerArgs(num_envs=n_envs, num_steps=T)
        self.learner = PPOLearner(self.agent, self.largs, world_size=world.world_size, chunk_rows=chunk_rows,
                                  use_graphs=use_graphs)
        data = make_rollout(T, n_envs, FRAMES, self.D, seed=1234, rank=world.rank)
        self.host = {k: v.pin_memory() for k, v in data.items()}
        self.dbuf = {k: v.to(dev, non_blocking=True) for k, v in self.host.items()}
        with torch.no_grad():       # what the rollout phase would have stored (ppo_atari_oc.py:505-509)
            action, logp, _, value = self.agent.get_action_and_value(self.dbuf["obs"].reshape(T * n_envs, FRAMES, self.D))
        self.dbuf["actions"], self.dbuf["logprobs"], self.dbuf["values"] = (action.reshape(T, n_envs), logp.reshape(T, n_envs),
                                                                           value.reshape(T, n_envs))
        for k in ("actions", "logprobs", "values"):
            self.host[k] = self.dbuf[k].cpu().pin_memory()
        # second set of persistent device buffers for the end-to-end steps (filled from the pinned host copies every step)
        self.ebuf = None
        torch.cuda.synchronize(dev)

    def step_device(self, perms):
        d, L = self.dbuf, self.learner
        adv, ret = L.compute_gae(d["rewards"], d["values"], d["dones"], d["next_obs"], d["next_done"])
        return L.update(d["obs"], d["actions"], d["logprobs"], adv, ret, d["values"], permutations=perms)

    H2D_KEYS = ("obs", "actions", "logprobs", "rewards", "dones", "values", "next_obs", "next_done")

    def step_e2e(self):
        if self.ebuf is None:
            self.ebuf = {k: torch.empty_like(self.dbuf[k]) for k in self.H2D_KEYS}
        d, L = self.ebuf, self.learner
        for k in self.H2D_KEYS:
            d[k].copy_(self.host[k], non_blocking=True)
        adv, ret = L.compute_gae(d["rewards"], d["values"], d["dones"], d["next_obs"], d["next_done"])
        # update() shuffles on the host as the reference does (:558-561), copies each epoch's order H2D, and reads the
        # statistics back (D2H) before returning
        return L.update(d["obs"], d["actions"], d["logprobs"], adv, ret, d["values"])

    def h2d_bytes(self):
        return (sum(self.host[k].numel() * self.host[k].element_size() for k in self.H2D_KEYS) +
                self.largs.update_epochs * self.T * self.N * 8)

    def d2h_bytes(self):
        from oc_cleanrl_b200 import _abi
        return self.largs.update_epochs * self.largs.num_minibatches * (_abi.OCPPO_NUM_STATS + 1) * 4

    def barrier(self):
        if self.world.distributed:
            torch.distributed.barrier()
        torch.cuda.synchronize(self.dev)

    def max_over_ranks(self, ms):
        if not self.world.distributed:
            return ms
        t = torch.tensor([ms], dtype=torch.float64, device=self.dev)
        torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
        return float(t.item())

    def timed(self, fn, steps, warmup):
        """W untimed + K timed calls of fn(i), barrier + synchronize on both sides, CUDA events, max over ranks."""
        for i in range(warmup):
            fn(i)
        self.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        last = None
        for i in range(steps):
            last = fn(warmup + i)
        e1.record()
        self.barrier()
        return self.max_over_ranks(e0.elapsed_time(e1)) / steps, last

    def samples_per_s(self, ms):
        return self.world.world_size * self.T * self.N / (ms * 1e-3)

    def free(self):
        self.agent = self.learner = self.dbuf = self.host = self.ebuf = None
        torch.cuda.empty_cache()


def run_b200(a):
    from oc_cleanrl_b200 import _abi
    from oc_cleanrl_b200 import distributed as ocd
    from oc_cleanrl_b200.learner import make_permutations
    from oc_cleanrl_b200.synthetic import GAME_SHAPES

    world = ocd.world_from_env()
    if not torch.cuda.is_available():
        raise SystemExit("bench.py (impl b200) needs a CUDA device; there is no CPU fallback")
    dev = torch.device("cuda", world.local_rank)
    torch.cuda.set_device(dev)
    ocd.init_process_group(world, "nccl")
    lib = _abi.lib()
    assert lib.ocppo_set_gemm_mode(ENGINES[a.gemm]) == 0
    D, A = GAME_SHAPES[a.game]
    T = a.num_steps
    strong_main = a.scaling == "strong"
    N = a.num_envs // world.world_size if strong_main else a.num_envs
    pk = peaks()
    use_graphs = not a.no_graphs
    rank_seed = ocd.seed_rank(1, world.rank)
    total = a.warmup + a.steps

    w = Workload(a, world, dev, a.game, N, T, a.enc, a.dec, a.chunk_rows, use_graphs, rank_seed)
    perms = [make_permutations(T * N, EPOCHS) for _ in range(total)]   # resident inputs of `value`

    # ---- `value`: device-resident inputs, profiling off ---------------------------------------
    sampler = ClockSampler(world.local_rank)
    if world.rank == 0:
        sampler.start()
    ms_per_step, last = w.timed(lambda i: w.step_device(perms[i]), a.steps, a.warmup)
    clocks = sampler.stop() if world.rank == 0 else None
    value_sps = w.samples_per_s(ms_per_step)
    graphs_live = bool(w.learner.use_graphs)
    graph_hits, graph_misses = w.learner._graph_hits, w.learner._graph_misses

    # ---- e2e: host buffers in, stats out, through the public API --------------------------------
    e2e = None
    if not a.no_e2e:
        ms_e2e, _ = w.timed(lambda i: w.step_e2e(), a.steps, 2)
        e2e = {"value": w.samples_per_s(ms_e2e), "unit": "samples/s", "ms_per_step": ms_e2e,
               "h2d_bytes_per_step": int(w.h2d_bytes()), "d2h_bytes_per_step": int(w.d2h_bytes())}

    # ---- per-class kernel time: a separate pass, plain launches, every launch bracketed by events --------
    w.learner.use_graphs = False
    w.step_device(perms[0])
    w.barrier()
    counts0, _ = _abi.launch_counts()
    lib.ocppo_profile_enable(1)
    p0, p1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    p0.record()
    for i in range(a.prof_steps):
        w.step_device(perms[i % total])
    p1.record()
    w.barrier()
    prof_ms, prof_n = _abi.profile_read()
    lib.ocppo_profile_enable(0)
    counts1, _ = _abi.launch_counts()
    prof_step_ms = p0.elapsed_time(p1) / a.prof_steps
    launched = {k: counts1[k] - counts0[k] for k in counts1}
    if sum(prof_n.values()) != sum(launched.values()) or any(prof_n[k] != launched[k] for k in launched):
        raise SystemExit(f"profiling pass dropped launches: timed {prof_n} vs launched {launched}")
    kernels_per_step = sum(launched.values()) / a.prof_steps
    w.learner.use_graphs = graphs_live

    # ---- strong scaling (N > 1): the same number of envs in TOTAL (ppo_atari_multigpu.py:166-172) ---------
    strong = None
    if world.distributed and not a.no_strong and not strong_main and a.num_envs % world.world_size == 0:
        ws = Workload(a, world, dev, a.game, a.num_envs // world.world_size, T, a.enc, a.dec, a.chunk_rows, use_graphs, rank_seed)
        sperms = [make_permutations(T * ws.N, EPOCHS) for _ in range(total)]
        ms_s, last_s = ws.timed(lambda i: ws.step_device(sperms[i]), a.steps, a.warmup)
        strong = {"scaling": "strong", "total_envs": a.num_envs, "num_envs_per_gpu": ws.N, "ms_per_step": ms_s,
                  "value": ws.samples_per_s(ms_s), "unit": "samples/s", "cuda_graphs": bool(ws.learner.use_graphs),
                  "loss": last_s["loss"],
                  "note": "same metric with the envs of ONE weak-scaling GPU split over all ranks; compare with the 1-GPU line's value"}
        if not a.no_e2e:
            ms_se, _ = ws.timed(lambda i: ws.step_e2e(), a.steps, 2)
            strong["e2e"] = {"value": ws.samples_per_s(ms_se), "unit": "samples/s", "ms_per_step": ms_se}
        ws.free()

    if world.rank != 0:
        if world.distributed:
            torch.distributed.destroy_process_group()
        return

    # ---- rooflines (rank 0) ---------------------------------------------------------
    tc = a.gemm != "fma"
    gemm_ms = (prof_ms["gemm_fwd"] + prof_ms["gemm_dgrad"] + prof_ms["gemm_wgrad"] + prof_ms["gemm_cuda_core"]) / a.prof_steps
    tc_ms = (prof_ms["gemm_fwd"] + prof_ms["gemm_dgrad"] + prof_ms["gemm_wgrad"]) / a.prof_steps
    gemm_flops = gemm_flops_per_step(D, N, T, a.enc, a.dec)
    gemm_tflops = gemm_flops / (gemm_ms * 1e-3) / 1e12 if gemm_ms > 0 else 0.0
    traffic = NCU_TRAFFIC.get((a.game, N, T, a.gemm)) if (a.enc, a.dec) == (ENC, DEC) else None
    roofline = {"bound": "tensor", "achieved": gemm_tflops, "peak": pk["bf16_sustained"], "unit": "TFLOP/s",
                "frac": gemm_tflops / pk["bf16_sustained"],
                "traffic": traffic["bytes"] if traffic else None, "traffic_source": traffic["source"] if traffic else None,
                "kernel": ("tc_gemm_kernel<FWD|DGRAD|WGRAD> (tcgen05.mma cta_group::2: " + ENGINE_TEXT.get(a.gemm, "") +
                           ", TMA-fed, TMEM accumulators, bulk-tensor-store epilogue) + the CUDA-core first-layer / skinny kernels"
                           if tc else "sgemm_kernel (fp32 CUDA-core FMA; fwd+dgrad+wgrad launches of the trunk)"),
                "peak_source": pk["source"] + ", bf16 dense sustained",
                "gemm_ms_per_step": gemm_ms, "tensor_core_kernel_ms_per_step": tc_ms,
                "share_of_step": gemm_ms / prof_step_ms,
                "launches_per_step": (prof_n["gemm_fwd"] + prof_n["gemm_dgrad"] + prof_n["gemm_wgrad"] + prof_n["gemm_cuda_core"]) / a.prof_steps,
                "flops_per_step": gemm_flops,
                "achieved_is": "ALGORITHMIC fp32-equivalent FLOPs (2*M*N*K per GEMM, SURVEY 8d) / CUDA-event time of the trunk GEMM "
                               "launches, from the profiled pass (plain launches, all of them timed)"}
    if tc:
        units = PASS_UNITS[a.gemm]
        # bf16-equivalent MMA work issued per algorithmic MAC: a TF32 pass costs 2 (half the bf16 rate), a BF16 / FP16 pass 1
        roofline["tensor_pipe"] = {"bf16_equivalent_tflops_issued": units * gemm_tflops, "peak": pk["bf16_sustained"],
                                   "frac_of_peak": units * gemm_tflops / pk["bf16_sustained"],
                                   "note": f"fp32-grade accuracy costs {units:.0f} bf16-equivalent MACs per algorithmic MAC, so frac = "
                                           f"achieved/peak is bounded by 1/{units:.0f} = {1.0 / units:.3f} at full tensor-pipe occupancy"}
    kernel_ms = {k: v / a.prof_steps for k, v in prof_ms.items() if v > 0}
    profile_pass = {"steps": a.prof_steps, "ms_per_step": prof_step_ms, "sum_kernel_ms_per_step": sum(kernel_ms.values()),
                    "kernels_timed": int(sum(prof_n.values())), "kernels_launched": int(sum(launched.values())),
                    "kernels_per_step_by_class": {k: v / a.prof_steps for k, v in launched.items() if v},
                    "note": "plain launches with a CUDA-event pair around each (adds ~2 us per launch); `value` is timed without"}

    sweep_n = [int(x) for x in a.gae_sweep.split(",") if x]
    sweep = []
    for n in sweep_n:
        us, us_copy, us_pdl = time_gae_isolated(dev, T, n)
        gbs = gae_bytes(T, n) / (us * 1e-6) / 1e9
        sweep.append({"num_envs": n, "num_steps": T, "us_per_launch": us, "bytes": gae_bytes(T, n), "GB/s": gbs,
                      "frac": gbs / pk["hbm"], "same_bytes_copy_us": us_copy,
                      "same_bytes_copy_frac": gae_bytes(T, n) / (us_copy * 1e-6) / 1e9 / pk["hbm"],
                      "overlapped_us_per_launch": us_pdl, "overlapped_frac": gae_bytes(T, n) / (us_pdl * 1e-6) / 1e9 / pk["hbm"]})
    own = next((s for s in sweep if s["num_envs"] == N), sweep[0] if sweep else None)
    roofline_gae = None
    if own:
        tg = NCU_TRAFFIC_GAE.get((own["num_envs"], T))
        roofline_gae = {"bound": "hbm", "achieved": own["GB/s"], "peak": pk["hbm"], "unit": "GB/s", "frac": own["frac"],
                        "traffic": tg["bytes"] if tg else None, "traffic_source": tg["source"] if tg else None,
                        "overlapped": {"us": own["overlapped_us_per_launch"], "frac": own["overlapped_frac"],
                                       "what": "the same launches as programmatic dependents of each other (ocppo_gae_overlapped, how the "
                                               "learner launches the scan behind the bootstrap value kernel): launch latency and the "
                                               "first DRAM round trip overlap the previous kernel's tail"},
                        "floor": {"us": own["same_bytes_copy_us"], "frac": own["same_bytes_copy_frac"],
                                  "what": "cudaMemcpyAsync D2D moving the same bytes, timed the same way (graph replay over >L2 of "
                                          "rotating buffers): launch + one DRAM round trip"},
                        "kernel": "gae_scan_kernel", "peak_source": pk["source"], "sweep": sweep}
    # rollout inference over the same sweep (BASELINE configs[4]); the trunk is 11.55 MFLOP per observation against
    # 116 bytes in and out, so it is held to the tensor roofline like the update, not to HBM
    inference = []
    units = PASS_UNITS[a.gemm] if tc else None
    for n in sweep_n:
        ms_api, ms_graph = time_policy_step(w.agent, dev, D, n)
        tf = 2.0 * trunk_macs(D, a.enc, a.dec) * n / (ms_graph * 1e-3) / 1e12
        inference.append({"num_envs": n, "ms_per_call": ms_graph, "ms_per_call_api": ms_api, "obs_per_s": n / (ms_graph * 1e-3),
                          "TFLOP/s": tf, "frac_of_bf16_peak": tf / pk["bf16_sustained"],
                          "tensor_pipe_frac": units * tf / pk["bf16_sustained"] if units else None,
                          "note": "ms_per_call: policy_step_into(graph=True), the rollout collector's route; "
                                  "ms_per_call_api: get_action_and_value (fresh outputs, plain launches)"})
    workspace = {"chunk_rows": int(w.agent.workspace_rows), "workspace_bytes": int(w.agent.workspace_bytes)}
    w.free()

    # ---- BASELINE configs[0] and configs[2] (1 GPU only) ---------------------------------------
    configs = None
    if world.world_size == 1 and not a.no_configs:
        configs = []
        for name, game, n_envs in (("configs[2] spaceinvaders", "spaceinvaders", 4096), ("configs[2] breakout", "breakout", 4096),
                                   ("configs[0] pong 8 envs", "pong", 8)):
            c = Workload(a, world, dev, game, n_envs, T, a.enc, a.dec, a.chunk_rows, use_graphs, rank_seed)
            cperms = [make_permutations(T * n_envs, EPOCHS) for _ in range(6)]
            ms_c, last_c = c.timed(lambda i: c.step_device(cperms[i]), 3, 3)
            rec = {"name": name, "game": game, "num_envs": n_envs, "num_steps": T, "obs_dim": c.D, "ms_per_step": ms_c,
                   "value": c.samples_per_s(ms_c), "unit": "samples/s", "steps": 3, "warmup": 3, "cuda_graphs": bool(c.learner.use_graphs),
                   "gemm_tflops_fp32_equivalent": gemm_flops_per_step(c.D, n_envs, T, a.enc, a.dec) / (ms_c * 1e-3) / 1e12,
                   "loss": last_c["loss"], "grad_norm": last_c["grad_norm"]}
            ms_ce, _ = c.timed(lambda i: c.step_e2e(), 3, 1)
            rec["e2e"] = {"value": c.samples_per_s(ms_ce), "unit": "samples/s", "ms_per_step": ms_ce}
            ms_api, ms_graph = time_policy_step(c.agent, dev, c.D, n_envs)
            rec["policy_step_ms"] = {"graph": ms_graph, "api": ms_api}
            configs.append(rec)
            c.free()

    cpu_baseline = None
    if world.world_size == 1 and not a.no_cpu_baseline:
        cpu = CpuReference(D, A, N, T, a.enc, a.dec)
        n_mb = max(1, a.ref_minibatches)
        cpu.gae(); cpu.minibatches(1)                                   # warm-up
        wall, sps, t_gae, t_upd = cpu.step(n_mb)
        cpu_baseline = {"value": sps, "unit": "samples/s", "cores": cpu.cores, "kind": "port", "sample": cpu.describe(n_mb),
                        "detail": {"torch_threads": torch.get_num_threads(), "gae_ms_full": 1e3 * t_gae,
                                   "s_per_true_size_minibatch": t_upd / n_mb, "wall_s": wall,
                                   "extrapolation_factor": EPOCHS * MINIBATCHES / n_mb}}

    knobs = {k: os.environ.get(k) for k in ("OCPPO_TC_DBG", "OCPPO_TC_OPS", "OCPPO_TC_TRUNC", "OCPPO_NO_MAPCACHE")}
    n_mb_step = EPOCHS * MINIBATCHES
    line = {
        "metric": "PPO learner samples/sec (GAE+update)", "value": value_sps, "unit": "samples/s",
        "n_gpus": world.world_size, "steps": a.steps, "warmup": a.warmup, "ms_per_step": ms_per_step,
        "higher_is_better": True, "scaling": a.scaling, "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(a, world.world_size, N, dict(
            workspace, cuda_graphs=graphs_live, library_version=int(lib.ocppo_version()),
            bringup_build=bool(lib.ocppo_bringup_build()),
            bringup_env=knobs if lib.ocppo_bringup_build() else "knobs compiled out (release build)")),
        "roofline": roofline, "roofline_gae": roofline_gae, "profile_pass": profile_pass,
        "inference": inference, "kernel_ms_per_step": kernel_ms, "cpu_baseline": cpu_baseline, "e2e": e2e,
        # kernels of this library EXECUTED inside the timed region: the per-step kernel count of the plain-launch pass
        # (+ one scalar-refresh kernel per minibatch in graph mode) x K; with CUDA graphs most of them are graph nodes
        "gpu_launches": int(round((kernels_per_step + (n_mb_step if graphs_live else 0)) * a.steps)),
        "graph_replays": int(n_mb_step * a.steps) if graphs_live else 0,
        "graph_cache": {"hits": graph_hits, "misses": graph_misses},
        "clocks": clocks,
        "last_stats": {k: last[k] for k in ("loss", "pg_loss", "v_loss", "entropy", "approx_kl", "grad_norm")},
        "impl": "b200",
    }
    if configs is not None:
        line["configs"] = configs
    if strong is not None:
        line["strong"] = strong
    emit(line)
    if world.distributed:
        torch.distributed.destroy_process_group()


if __name__ == "__main__":
    args = parse()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)
