#!/usr/bin/env python
"""bench.py — env-steps/sec of the PPO hot path (rollout + GAE + update), device-timed.

    python bench.py --gpus N --steps K --warmup W            # this repo's sm_100a path
    python bench.py --impl reference --steps K --warmup W    # the reference's own CPU path (oracle/_ref/ppo_ref)

A "step" is one PPO iteration of BASELINE.json configs[4] (the north-star shape): CartPole PPO_Discrete,
1 048 576 envs x 128 steps IN TOTAL, 2x64 MLP, E=10 epochs x 4 minibatches (CartPoleRecommendedSettings.toml
hyper-parameters, action_size=2).  N>1 (torchrun, one rank per GPU): the envs shard by rank (STRONG scaling,
1 048 576 / N envs per GPU); per optimiser step the gradient is exchanged over NVLink peer memory, the advantage
moments go through one NCCL allreduce per iteration.  The latency-bound point, configs[1] (4096 envs x 128 steps per
GPU, weak), is measured in the same run and reported under "config_b".  Rank 0 prints ONE JSON line.
"""
import argparse
import json
import os
import re
import shutil
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

WORKLOADS = {
    # name: env, O, nvec, H, envs per GPU, steps, minibatches, epochs, max_episode_steps
    "cartpole_4096x128_h64": dict(env=0, O=4, nvec=(2,), H=64, N=4096, T=128, MB=4, E=10, max_ep=500),
    "cartpole_131072x128_h64": dict(env=0, O=4, nvec=(2,), H=64, N=131072, T=128, MB=4, E=10, max_ep=500),
    # total=True: N is the WHOLE job's env count, sharded N / world per GPU (strong scaling)
    "cartpole_1048576x128_h64": dict(env=0, O=4, nvec=(2,), H=64, N=1048576, T=128, MB=4, E=10, max_ep=500, total=True),
    "mountaincar_65536x256_h256": dict(env=1, O=2, nvec=(3,), H=256, N=65536, T=256, MB=4, E=10, max_ep=200),
    "cartpole_8x32_h64": dict(env=0, O=4, nvec=(2,), H=64, N=8, T=32, MB=4, E=10, max_ep=500),
    # BASELINE configs[3] through the trainer: PPO_MultiDiscrete, 8 CategoricalMasked heads (nvec {3,3,3,3,3,2,2,2}), 262 144 envs of
    # the synthetic MultiWalk environment (state-dependent masks; ppo-libtorch_b200/csrc/env.cuh)
    "multiwalk_262144x32_h64": dict(env=2, O=4, nvec=(3, 3, 3, 3, 3, 2, 2, 2), H=64, N=262144, T=32, MB=4, E=10, max_ep=50,
                                    masked=True),
}
ENV_NAMES = {0: "CartPole", 1: "MountainCar", 2: "MultiWalk (synthetic, masked MultiDiscrete)"}
DEFAULT_WORKLOAD = "cartpole_1048576x128_h64"
CONFIG_B = "cartpole_4096x128_h64"
REF_SAMPLE_ENVS = 32768   # env count the CPU arm runs when the workload itself would take hours on host cores
# dram__bytes_read.sum + dram__bytes_write.sum of ONE update-kernel launch from an `ncu --set full` capture of this
# command (profiles/r01r_update_tc_ncu.txt, cold L2): 14.82 MB.  Algorithmic bytes per launch are 131072 rows x
# 36 B = 4.7 MB, fetched as 64-byte packed records = 8.4 MB (ppo_pack_rows), + 37 KB of parameters per CTA.
NCU_TRAFFIC = {("cartpole_4096x128_h64", 2): 14.82e6,   # (capture of the round-1 kernel)
               # north-star shape on one GPU, 33.5 M rows per launch (profiles/r02f_update_tc_default_ncu.txt): 8.14 GB read + 11 MB
               # written = 243 B per row.  Algorithmic: 36 B per row, fetched as a 64-byte packed record by BOTH warpgroups of a
               # tile stream (actor and critic each read the row; free-running they drift ~15 tiles apart, so the second read
               # misses L2).  Keeping the critic within 1-3 tiles of the actor halves this to 4.28 GB (measured) but costs 1.5-1.8 %
               # of the step (A/B on one box), and DRAM is at 16 % of its peak here: not done.
               ("cartpole_1048576x128_h64", 2): 8.154e9,
               # hidden = 256 (profiles/r02a_update_h256_ncu.txt): update_h256_kernel 1.04 GB read + 8.55 GB written (the fp16 h1 /
               # dZ2 scratch of 4.19 M rows x 2 nets), dw2_h256_kernel 8.60 GB read + 0.07 GB written
               ("mountaincar_65536x256_h256", 2): 18.25e9}
HYPER = dict(gamma=0.98, gae_lambda=0.95, clip_coef=0.2, norm_adv=True, clip_vloss=True, ent_coef=0.0, vf_coef=0.5,
             max_grad_norm=0.5)
LR0 = 1e-3


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return dict(hbm=float(d["hbm_gbs"]), bf16=float(d["bf16_tflops"]), bf16_sustained=float(
            d.get("bf16_tflops_sustained", d["bf16_tflops"])), source="measured")
    return dict(hbm=6650.0, bf16=1590.0, bf16_sustained=1400.0, source="fallback")


def fwd_flops(O, H, K):
    """FLOPs of one actor + critic forward per sample"""
    return 2 * ((O * H + H * H + H * 1) + (O * H + H * H + H * K))


# ------------------------------------------------------------------------------------------- clocks
class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)"""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.proc, self.lines, self.gpu = None, [], gpu_index

    def start(self):
        if not shutil.which("nvidia-smi"):
            return
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", str(self.gpu)], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1]))
                mx.append(float(f[2]))
            except ValueError:
                continue
            for nm, val in zip(names, f[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(nm)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ------------------------------------------------------------------------------------------- reference arm
def run_reference(args, wl_name):
    """The reference's own CPU implementation (PPO_Discrete::train: ThreadPool envs + LibTorch CPU update), the
    unmodified sources compiled by oracle/Makefile into oracle/_ref/ppo_ref, run on this box's host cores with
    a PPOConfig.toml for the same workload.  A step = one PPO update of the reference; the sample is bounded so
    the whole run ends within a few minutes."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    wl = dict(WORKLOADS[wl_name])
    binary = os.path.join(ROOT, "oracle", "_ref", "ppo_ref")
    cores = os.cpu_count() or 1
    total = bool(wl.get("total"))
    base = {"impl": "reference", "metric": "env_steps_per_sec", "unit": "env-steps/s", "n_gpus": args.gpus,
            "higher_is_better": True, "scaling": "strong" if total else "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic",
            "config": {"workload": wl_name, "envs": wl["N"], "rollout_steps": wl["T"], "hidden": wl["H"],
                       "update_epochs": wl["E"], "minibatches": wl["MB"]}}
    ref_envs = wl["N"]
    if total and wl["N"] > REF_SAMPLE_ENVS:
        # 1 048 576 envs x 128 steps = 134 M env-steps per update: ~15 min per update on host cores.  The CPU arm runs
        # the same PPO update on a REF_SAMPLE_ENVS-env shard of it (its env-steps/s is flat in the env count from a few
        # thousand envs up: per-env ThreadPool jobs and per-sample LibTorch work both scale linearly) and the rate is
        # quoted for the workload as is.
        ref_envs = REF_SAMPLE_ENVS
        wl["N"] = ref_envs
        base["config"]["reference_sample_envs"] = ref_envs
    if wl["env"] == 2:
        print(json.dumps(dict(base, unavailable="MultiWalk is a synthetic environment of this repo: the reference has no CPU implementation of it")))
        return 0
    if not os.path.exists(binary):
        print(json.dumps(dict(base, unavailable="oracle/_ref/ppo_ref not built (needs /root/reference at build time)")))
        return 0
    batch = wl["N"] * wl["T"]
    # bound the sample: the reference does ~5e4 env-steps/s on 8 cores at this size; cap the run near 150 s
    want = args.steps + args.warmup
    est_rate = 5.0e4 * max(1.0, cores / 8.0) ** 0.5
    cap = max(2, int(150.0 * est_rate / batch))
    updates = min(want, cap)
    warm = min(args.warmup, updates - 1)
    tmp = tempfile.mkdtemp(prefix="ppo_ref_")
    toml = f"""[environment]
obs_size = {wl['O']}
action_size = {wl['nvec'][0]}
max_episode_steps = {wl['max_ep']}
[general]
seed = 2
total_timesteps = {updates * batch}
use_cuda = false
torch_deterministic = true
checkpoint_updates = 1000000
[ppo]
learning_rate = {LR0}
num_envs = {wl['N']}
num_steps = {wl['T']}
anneal_lr = true
use_gae = true
gamma = {HYPER['gamma']}
gae_lambda = {HYPER['gae_lambda']}
num_minibatches = {wl['MB']}
update_epochs = {wl['E']}
norm_adv = true
clip_coef = {HYPER['clip_coef']}
clip_vloss = true
ent_coef = {HYPER['ent_coef']}
vf_coef = {HYPER['vf_coef']}
max_grad_norm = {HYPER['max_grad_norm']}
"""
    with open(os.path.join(tmp, "PPOConfig.toml"), "w") as f:
        f.write(toml)
    t0 = time.time()
    # torchrun exports OMP_NUM_THREADS=1 to its workers: the LibTorch CPU update must see every host core
    env = {k: v for k, v in os.environ.items() if k not in ("OMP_NUM_THREADS", "MKL_NUM_THREADS", "OMP_PROC_BIND")}
    try:
        out = subprocess.run([binary], cwd=tmp, capture_output=True, text=True, timeout=1500, env=env).stdout
    except subprocess.TimeoutExpired as e:
        out = (e.stdout or b"").decode() if isinstance(e.stdout, bytes) else (e.stdout or "")
    wall = time.time() - t0
    shutil.rmtree(tmp, ignore_errors=True)
    fps = [float(x) for x in re.findall(r"\|\s+fps\s+\|\s+(\d+)", out)]
    if not fps:
        print(json.dumps(dict(base, unavailable="reference binary printed no fps rows: " + out[-200:].replace("\n", " "))))
        return 0
    timed = fps[warm:] if len(fps) > warm else fps
    ms = [batch / f * 1000.0 for f in timed]
    ms_per_step = sum(ms) / len(ms)
    value = batch / (ms_per_step / 1000.0)
    sample = (f"{len(timed)} PPO update(s) of {batch} env-steps each ({ref_envs} envs x {wl['T']} steps) after {warm} "
              f"warm-up update(s) (requested {args.steps}+{args.warmup}, bounded to {updates}); self-reported "
              f"per-update fps of the unmodified reference binary on {cores} host cores (OMP/MKL thread caps removed "
              f"from its environment); wall {wall:.1f} s")
    line = dict(base, value=value, steps=len(timed), warmup=warm, ms_per_step=ms_per_step,
                cpu_baseline={"value": value, "unit": "env-steps/s", "cores": cores, "kind": "reference",
                              "sample": sample},
                e2e={"value": value, "unit": "env-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                gpu_launches=0)
    print(json.dumps(line))
    return 0


def cpu_baseline_sample(wl_name):
    """cpu_baseline leg of the b200 arm: the reference binary on a bounded sample (2 updates, first discarded)"""
    class A:
        steps, warmup, gpus = 1, 1, 1
    import io
    from contextlib import redirect_stdout
    buf = io.StringIO()
    with redirect_stdout(buf):
        run_reference(A, wl_name)
    try:
        d = json.loads(buf.getvalue().strip().splitlines()[-1])
        return d.get("cpu_baseline") or {"value": None, "unit": "env-steps/s", "cores": os.cpu_count(),
                                         "kind": "reference", "sample": d.get("unavailable", "unavailable")}
    except Exception as e:   # noqa: BLE001
        return {"value": None, "unit": "env-steps/s", "cores": os.cpu_count(), "kind": "reference",
                "sample": f"failed: {e}"}


# ------------------------------------------------------------------------------------------- kernel rooflines
def kernel_rooflines(P, torch, pk):
    """HBM-bound kernels timed alone at their graded sizes, CUDA events, L2 flushed by size
    (inputs >> 126 MB L2) — achieved = algorithmic bytes / mean launch time."""
    out = {}
    dev = "cuda"

    def timeit(fn, iters=10):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        for _ in range(iters):
            fn()
        e.record()
        torch.cuda.synchronize()
        return s.elapsed_time(e) / iters

    # GAE at config E's per-GPU shard [128, 131072] x 8 (= full 1M envs on one GPU) -> 2.68 GB of traffic
    T, N = 128, 1048576
    r = torch.randn(T, N, device=dev)
    v = torch.randn(T, N, device=dev)
    d = (torch.rand(T, N, device=dev) < 0.05).float()
    nv = torch.randn(N, device=dev)
    nd = torch.zeros(N, device=dev)
    adv, ret = torch.empty_like(r), torch.empty_like(r)
    ms = timeit(lambda: P.ops.gae(r, v, d, nv, nd, 0.98, 0.95, out=(adv, ret)))
    by = 20.0 * T * N
    out["gae"] = {"shape": [T, N], "ms": ms, "bytes": by, "achieved": by / ms / 1e6, "peak": pk["hbm"],
                  "unit": "GB/s", "frac": by / ms / 1e6 / pk["hbm"]}
    del r, v, d, adv, ret
    # env step, CartPole, 16M envs (60 B/env-step + 16 B obs row)
    n = 1 << 24
    state, ep_len, ep_ret, obs = P.ops.env_reset(P.ENV_CARTPOLE, n, seed=1)
    act = (torch.rand(n, device=dev) < 0.5).to(torch.int32)
    bufs = [torch.empty((n, 4), device=dev), torch.empty(n, device=dev), torch.empty(n, device=dev)]
    import ctypes as C
    lib = P.capi.load()

    def step():
        P.capi.check(lib.ppo_env_step(0, n, P.capi.ptr(state), P.capi.ptr(ep_len), P.capi.ptr(ep_ret),
                                      P.capi.ptr(act), 500, None, 1, 7, P.capi.ptr(bufs[0]), P.capi.ptr(bufs[1]),
                                      P.capi.ptr(bufs[2]), None, None, P.capi.stream_ptr()))
    ms = timeit(step)
    by = (60.0 + 16.0) * n      # 60 B/env-step (state r+w 32, action 4, reward 4, done 4, ep_len + ep_ret r+w 16) + 16 B obs row
    out["env_step_cartpole"] = {"envs": n, "ms": ms, "bytes": by, "achieved": by / ms / 1e6, "peak": pk["hbm"],
                                "unit": "GB/s", "frac": by / ms / 1e6 / pk["hbm"]}
    del state, ep_len, ep_ret, obs, act, bufs
    # env step, MountainCar, 16M envs (44 B/env-step: state r+w 16, action 4, reward 4, done 4, counters r+w 16; + 8 B obs row)
    state, ep_len, ep_ret, obs = P.ops.env_reset(P.ENV_MOUNTAINCAR, n, seed=1)
    act = torch.randint(0, 3, (n,), device=dev, dtype=torch.int32)
    bufs = [torch.empty((n, 2), device=dev), torch.empty(n, device=dev), torch.empty(n, device=dev)]

    def step_mc():
        P.capi.check(lib.ppo_env_step(1, n, P.capi.ptr(state), P.capi.ptr(ep_len), P.capi.ptr(ep_ret),
                                      P.capi.ptr(act), 200, None, 1, 7, P.capi.ptr(bufs[0]), P.capi.ptr(bufs[1]),
                                      P.capi.ptr(bufs[2]), None, None, P.capi.stream_ptr()))
    ms = timeit(step_mc)
    by = (44.0 + 8.0) * n
    out["env_step_mountaincar"] = {"envs": n, "ms": ms, "bytes": by, "achieved": by / ms / 1e6, "peak": pk["hbm"],
                                   "unit": "GB/s", "frac": by / ms / 1e6 / pk["hbm"]}
    del state, ep_len, ep_ret, obs, act, bufs
    # categorical, masked, nvec {3,3,3,3,3,2,2,2}, 4M rows (145 B/row, Philox uniforms)
    n, nvec = 1 << 22, [3, 3, 3, 3, 3, 2, 2, 2]
    logits = torch.randn(n, 21, device=dev)
    mask = (torch.rand(n, 21, device=dev) < 0.7).to(torch.uint8)
    mask[:, [0, 3, 6, 9, 12, 15, 17, 19]] = 1
    a = torch.empty((n, 8), dtype=torch.int32, device=dev)
    lp, en = torch.empty(n, device=dev), torch.empty(n, device=dev)
    nv_c = (C.c_int32 * 8)(*nvec)

    def cat():
        P.capi.check(lib.ppo_categorical(n, 8, nv_c, P.capi.ptr(logits), P.capi.ptr(mask), None, None, 1, 2, 0,
                                         P.capi.ptr(a), P.capi.ptr(lp), P.capi.ptr(en), None, P.capi.stream_ptr()))
    ms = timeit(cat)
    by = 145.0 * n
    out["categorical_masked_k21"] = {"rows": n, "ms": ms, "bytes": by, "achieved": by / ms / 1e6, "peak": pk["hbm"],
                                     "unit": "GB/s", "frac": by / ms / 1e6 / pk["hbm"]}
    # the same at BASELINE configs[3]'s own size (262 144 rows, 38 MB: L2-resident and launch-latency sized)
    full = n
    n = 262144
    ms = timeit(cat, iters=50)
    by = 145.0 * n
    out["categorical_masked_k21_config_d"] = {"rows": n, "ms": ms, "bytes": by, "achieved": by / ms / 1e6,
                                              "peak": pk["hbm"], "unit": "GB/s", "frac": by / ms / 1e6 / pk["hbm"],
                                              "note": "38 MB working set: L2-resident between repetitions"}
    n = full
    return out


# ------------------------------------------------------------------------------------------- b200 arm
def make_trainer(P, torch, dist, wl, n_local, rank, world, args):
    """one trainer per rank on its env shard; N>1: NCCL communicator (advantage moments) + peer-memory gradient exchange"""
    tr = P.Trainer(env_kind=wl["env"], obs_dim=wl["O"], hidden=wl["H"], nvec=wl["nvec"], num_envs=n_local,
                   num_steps=wl["T"], num_minibatches=wl["MB"], update_epochs=wl["E"], max_episode_steps=wl["max_ep"],
                   seed=2, world_size=world, rank=rank, use_cuda_graph=not args.no_graph, masked=bool(wl.get("masked")),
                   **HYPER)
    exchange = "none"
    if world > 1:
        uid = [P.unique_id() if rank == 0 else None]
        dist.broadcast_object_list(uid, src=0)
        tr.init_comm(uid[0], rank, world)
        exchange = "nccl"
        if not args.no_peer:
            # gradient exchange over peer memory (one fused kernel per rank) instead of reduce -> NCCL -> clip/AdamW
            hs = [None] * world
            dist.all_gather_object(hs, tr.peer_handle())
            try:
                tr.set_peers(hs)
                exchange = "peer-memory fused reduce+AdamW (NVLink P2P)"
            except P.PpoError as e:
                if rank == 0:
                    print("peer-memory path unavailable, NCCL path in use:", e, file=sys.stderr)
    return tr, exchange


def time_trainer(torch, dist, tr, K, W, flush, it0_total):
    """W warm-up iterations, then K device-timed iterations (L2 flushed, untimed, before each) twice: fully asynchronous
    (value) and through the public call with the host result (e2e).  Returns (ms_dev, ms_e2e, wall_dev), max over ranks."""
    def lr_at(i):
        return (1.0 - i / max(1, it0_total)) * LR0        # anneal_lr (PPO_Discrete.cpp:514-518)

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def timed_loop(want_result, it0):
        evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]
        barrier()
        for i in range(K):
            flush.zero_()
            evs[i][0].record()
            tr.iterate(lr_at(it0 + i), want_result=want_result)
            evs[i][1].record()
        barrier()
        return sum(s.elapsed_time(e) for s, e in evs)

    for i in range(W):
        tr.iterate(lr_at(i), want_result=False)
    barrier()
    t_wall = time.time()
    ms_dev = timed_loop(False, W)              # inputs resident in HBM, fully asynchronous iterations
    wall_dev = time.time() - t_wall
    ms_e2e = timed_loop(True, W + K)           # public call with host result: scalars H2D, diagnostics D2H, sync

    def max_over_ranks(x):
        if dist is None:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return t.item()
    return max_over_ranks(ms_dev), max_over_ranks(ms_e2e), wall_dev


UPDATE_PATHS = {0: "fp32 FMA kernels (update_kernel)",
                1: "tcgen05, hidden = 64, one fused kernel (update_tc_kernel: one net per compute warpgroup, 16 compute warps per SM)",
                2: "tcgen05, hidden = 256: update_h256_kernel (forward + loss + dX, one net per CTA) + dw2_h256_kernel (TMA-fed dW2 GEMM)"}


# what actually bounds the update kernel of each path (the dense-bf16 peak is the contract's denominator)
UPDATE_LIMITER = {
    1: "shared-memory bandwidth, not the tensor pipe: 64-wide layers make every SS-mode tcgen05.mma cost its operand bytes / "
       "128 B per clock (profiles/r02d_umma_conc.txt: M128N64K16 48 cycles for 32 cycles of math, M64N8K16 23), and the "
       "activation tiles cross shared memory 4 times per net and row (LSU wavefronts 54 % of peak + operand fetch ~33 %)",
    2: "tensor pipe / MUFU / issue slots in about equal thirds (update_h256_kernel); dw2_h256_kernel is at the HBM roofline"}


def run_b200(args, wl_name):
    import ctypes as C
    import torch
    import ppo_libtorch_b200 as P
    wl = WORKLOADS[wl_name]
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.gpus != world:
        if world == 1 and args.gpus > 1:
            sys.exit(f"--gpus {args.gpus} needs torchrun with {args.gpus} ranks (WORLD_SIZE is 1)")
    if not torch.cuda.is_available():
        sys.exit("bench.py: no CUDA device — the sm_100a path has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    pk = peaks()
    total = bool(wl.get("total"))
    n_local = wl["N"] // world if total else wl["N"]
    tr, exchange = make_trainer(P, torch, dist, wl, n_local, rank, world, args)

    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")   # > 126 MB L2
    K, W = args.steps, args.warmup
    clocks = ClockSampler(local)
    if rank == 0:
        clocks.start()
    ms_dev, ms_e2e, wall_dev = time_trainer(torch, dist, tr, K, W, flush, 2 * (K + W))
    clk = clocks.stop() if rank == 0 else None
    batch = n_local * wl["T"]
    total_steps = batch * world * K
    value = total_steps / (ms_dev / 1000.0)
    e2e_value = total_steps / (ms_e2e / 1000.0)

    # per-phase device time inside real iterations (cudaEvent pairs on the launch stream, eager launches)
    tr.profile(True)
    prof_iters = max(1, min(K, 3))
    for i in range(prof_iters):
        tr.iterate(LR0 * 0.5, want_result=False)
    phases = tr.phase_ms()
    tr.profile(False)
    n_opt = wl["E"] * wl["MB"]
    upd_ms, upd_cnt = phases["update"]
    ff = fwd_flops(wl["O"], wl["H"], sum(wl["nvec"]))
    mb_rows = batch // wl["MB"]
    flops_per_launch = 3.0 * ff * mb_rows        # forward + dX + dW  (update ~ 3 x forward)
    upd_ms_per = upd_ms / max(1, upd_cnt)
    achieved_tf = flops_per_launch / (upd_ms_per / 1000.0) / 1e12
    phase_total = sum(v[0] for v in phases.values())
    mode = P.ops.get_math_mode()
    path = int(P.capi.load().ppo_update_path(C.byref(tr.cfg.dims)))     # which kernels serve this shape in this mode
    traffic = NCU_TRAFFIC.get((wl_name, mode)) if (world == 1 or not total) else None   # per-launch captures are per shard size
    roofline = {"bound": "tensor", "kernel": "fused fwd + PPO loss + bwd of one minibatch: " + UPDATE_PATHS.get(path, "?"),
                "achieved": achieved_tf, "peak": pk["bf16_sustained"], "unit": "TFLOP/s",
                "frac": achieved_tf / pk["bf16_sustained"],
                "traffic": traffic,
                "traffic_source": None if traffic is None else
                "constant from the committed ncu --set full capture named beside NCU_TRAFFIC in bench.py, not re-measured by this run",
                "peak_source": pk["source"] + " (bf16 cuBLAS sustained)",
                "limiter": UPDATE_LIMITER.get(path),
                "flops_per_launch": flops_per_launch, "rows_per_launch": mb_rows, "ms_per_launch": upd_ms_per,
                "share_of_step": upd_ms / phase_total if phase_total else None}
    # one GPU / peer-memory DP: scalars, rollout, gae, all-epoch advantage statistics, row packing, per minibatch {update,
    # reduce+clip+AdamW}, explained variance.  NCCL DP: per epoch {permutation, moments, finalize}, per minibatch
    # {update, reduce, clip_adamw} (NCCL's own kernels are not counted: not ours).
    upd_launches = {0: 1, 1: 1, 2: 2}.get(path, 1)      # kernels per minibatch gradient
    if world == 1 or exchange.startswith("peer"):
        packed = 1 if (len(wl["nvec"]) == 1 and wl["O"] <= 4 and wl["H"] in (64, 256)) else 0   # + ppo_pack_rows
        launches = 4 + packed + 1 + wl["E"] * wl["MB"] * (upd_launches + 1) + 1 + (1 if world > 1 else 0)   # + episode tail
    else:
        launches = 3 + wl["E"] * (3 + wl["MB"] * (upd_launches + 2)) + 1

    line = {"metric": "env_steps_per_sec", "value": value, "unit": "env-steps/s", "n_gpus": world, "steps": K,
            "warmup": W, "ms_per_step": ms_dev / K, "higher_is_better": True, "scaling": "strong" if total else "weak",
            "vs_baseline": None,
            "dtype": {0: "f32", 1: "bf16", 2: "f16"}[mode], "data": "synthetic",
            "config": {"workload": wl_name, "env": ENV_NAMES[wl["env"]],
                       "envs_total": n_local * world, "envs_per_gpu": n_local, "rollout_steps": wl["T"],
                       "hidden": wl["H"], "update_epochs": wl["E"],
                       "minibatches": wl["MB"], "global_batch": batch * world, "parallelism": f"dp{world}",
                       "grad_exchange": exchange,
                       "l2": "flushed between steps (256 MiB write, untimed)", "cuda_graph": not args.no_graph,
                       "policy": "orthogonal-init weights, live-sampled actions"},
            "e2e": {"value": e2e_value, "unit": "env-steps/s", "h2d_bytes_per_step": 32,
                    "d2h_bytes_per_step": n_opt * 32 + 32, "ms_per_step": ms_e2e / K},
            "gpu_launches": launches * K, "roofline": roofline,
            "phases_ms_per_step": {k: v[0] / prof_iters for k, v in phases.items()},
            "clocks": clk, "wall_s_timed_region": wall_dev}
    tr.close()
    del tr
    if wl_name != CONFIG_B and not args.no_config_b:
        # the latency-bound point: BASELINE configs[1], 4096 envs x 128 steps PER GPU (weak), same run, same box
        wb = WORKLOADS[CONFIG_B]
        trb, _ = make_trainer(P, torch, dist, wb, wb["N"], rank, world, args)
        Kb = max(K, 20)
        mb_dev, mb_e2e, _ = time_trainer(torch, dist, trb, Kb, max(W, 5), flush, 4 * Kb)
        steps_b = wb["N"] * wb["T"] * world * Kb
        line["config_b"] = {"workload": CONFIG_B, "scaling": "weak", "envs_per_gpu": wb["N"],
                            "value": steps_b / (mb_dev / 1000.0), "ms_per_step": mb_dev / Kb,
                            "e2e_value": steps_b / (mb_e2e / 1000.0), "steps": Kb, "unit": "env-steps/s"}
        trb.close()
        del trb
    if rank == 0 and world == 1 and not args.no_kernels:
        line["kernels"] = kernel_rooflines(P, torch, pk)
    if rank == 0 and world == 1 and not args.no_cpu:
        line["cpu_baseline"] = cpu_baseline_sample(wl_name)
    elif rank == 0:
        line["cpu_baseline"] = {"value": None, "unit": "env-steps/s", "cores": os.cpu_count(), "kind": "reference",
                                "sample": "measured at N=1 only"}
    if rank == 0:
        print(json.dumps(line))
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default=DEFAULT_WORKLOAD, choices=sorted(WORKLOADS))
    ap.add_argument("--no-graph", action="store_true", help="eager launches instead of CUDA-graph replay")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-peer", action="store_true", help="N>1: NCCL gradient allreduce instead of the peer-memory kernel")
    ap.add_argument("--no-kernels", action="store_true", help="skip the stand-alone kernel rooflines")
    ap.add_argument("--no-config-b", action="store_true", help="skip the extra configs[1] (4096 envs per GPU) measurement")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup
    if args.impl == "reference":
        return run_reference(args, args.workload)
    return run_b200(args, args.workload)


if __name__ == "__main__":
    sys.exit(main())
