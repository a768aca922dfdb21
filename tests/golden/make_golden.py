"""Generate tests/golden/*.npz from the COMPILED REFERENCE (oracle/_ref/libppo_ref.so).

Run in the build container (where /root/reference exists):
    make -C oracle ref && python tests/golden/make_golden.py

Every array below is an output of the unmodified reference classes on seeded inputs
(see oracle/ref_shim.cpp for which reference symbol produces what).  The fixtures are what pins
oracle/ppo_oracle.c -- and through it the CUDA kernels -- on machines where the reference is absent.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import oracle  # noqa: E402

NVEC_MULTI = np.array([3, 3, 3, 3, 3, 2, 2, 2], np.int64)   # PPO/Agent.cpp:21 (commented nvec)


def save(name, **kw):
    np.savez_compressed(os.path.join(HERE, name + ".npz"), **kw)
    print("wrote", name, {k: getattr(v, "shape", None) for k, v in kw.items()})


def main():
    ref = oracle.ref()
    assert ref is not None, "build oracle/_ref first (make -C oracle ref)"
    rng = np.random.default_rng(20261104)

    # ---- CartPole: reset stream for seed 2 (shipped toml) and single steps
    resets = ref.cartpole_reset_seq(2, 8)
    n = 768
    st = rng.uniform(-1, 1, (n, 4)).astype(np.float32) * np.array([2.6, 3.0, 0.23, 3.0], np.float32)
    st[:64] = rng.uniform(-0.05, 0.05, (64, 4)).astype(np.float32)
    act = rng.integers(0, 2, n).astype(np.int64)
    out, rew, term = ref.cartpole_step(st, act)
    save("cartpole", resets_seed2=resets, state=st, action=act, next_state=out, reward=rew, terminated=term)

    # ---- MountainCar single steps (incl. walls and goal)
    st = np.stack([rng.uniform(-1.25, 0.65, n), rng.uniform(-0.075, 0.075, n)], 1).astype(np.float32)
    st[:16, 0] = -1.2
    st[16:32, 0] = 0.5
    st[32:48, 0] = 0.6
    act = rng.integers(0, 3, n).astype(np.int64)
    out, rew, term = ref.mountaincar_step(st, act)
    save("mountaincar", state=st, action=act, next_state=out, reward=rew, terminated=term)

    # ---- vectorised stepping with truncation + auto-reset (all envs seeded alike: quirk Q4)
    N, steps, seed, max_ep = 8, 160, 2, 30
    actions = rng.integers(0, 2, (steps, N)).astype(np.int64)
    obs, rew, done = ref.cartpole_vec_rollout(N, steps, seed, max_ep, actions)
    save("cartpole_vec", N=N, steps=steps, seed=seed, max_episode_steps=max_ep, actions=actions,
         obs=obs, reward=rew, done=done)

    # ---- distributions
    d = {}
    for K in (1, 2, 3, 7):
        lg = (rng.standard_normal((257, K)) * 2.0).astype(np.float32)
        a = rng.integers(0, K, 257).astype(np.int64)
        nl, pr, lp, ent = ref.categorical(lg, a)
        d.update({f"K{K}_logits": lg, f"K{K}_actions": a, f"K{K}_normlogits": nl, f"K{K}_probs": pr,
                  f"K{K}_logprob": lp, f"K{K}_entropy": ent})
    save("categorical", **d)
    d = {}
    for K in (3, 6, 21):
        lg = (rng.standard_normal((257, K)) * 2.0).astype(np.float32)
        mk = (rng.uniform(size=(257, K)) < 0.6).astype(np.uint8)
        mk[:, 0] = 1
        # sample valid actions
        a = np.array([rng.choice(np.flatnonzero(m)) for m in mk], np.int64)
        nl, pr, lp, ent = ref.categorical(lg, a, mask=mk)
        d.update({f"K{K}_logits": lg, f"K{K}_mask": mk, f"K{K}_actions": a, f"K{K}_normlogits": nl,
                  f"K{K}_probs": pr, f"K{K}_logprob": lp, f"K{K}_entropy": ent})
    save("categorical_masked", **d)

    # ---- agent forward / evaluate (orthogonal init from the reference itself)
    d = {}
    for tag, O, K in (("cartpole", 4, 2), ("cartpole_k1", 4, 1), ("mountaincar", 2, 3)):
        tr = ref.trainer(O, K, lr=1e-3, seed=1)
        params = tr.get_params()
        obs = rng.standard_normal((96, O)).astype(np.float32)
        logits, value = tr.forward(obs)
        a = rng.integers(0, K, 96).astype(np.int64)
        lp, ent, val = tr.evaluate(obs, a)
        d.update({f"{tag}_params": params, f"{tag}_obs": obs, f"{tag}_logits": logits, f"{tag}_value": value,
                  f"{tag}_actions": a, f"{tag}_logprob": lp, f"{tag}_entropy": ent})
        tr.close()
    # masked multi-head: nvec {3,3,3,3,3,2,2,2}
    tr = ref.trainer(2, int(NVEC_MULTI.sum()), lr=1e-3, seed=3)
    tr.set_nvec(NVEC_MULTI)
    # the init gives near-uniform heads; scale the actor head so masks/logits matter
    params = tr.get_params()
    params[-(21 * 64 + 21):] *= 60.0
    tr.set_params(params)
    obs = rng.standard_normal((96, 2)).astype(np.float32)
    mk = (rng.uniform(size=(96, 21)) < 0.6).astype(np.uint8)
    offs = np.concatenate([[0], np.cumsum(NVEC_MULTI)[:-1]])
    mk[:, offs] = 1
    acts = np.zeros((96, 8), np.int64)
    for r in range(96):
        for h, (o, k) in enumerate(zip(offs, NVEC_MULTI)):
            acts[r, h] = rng.choice(np.flatnonzero(mk[r, o:o + k]))
    lp, ent, val = tr.evaluate(obs, acts, mask=mk)
    logits, value = tr.forward(obs)
    d.update(multi_params=params, multi_obs=obs, multi_mask=mk, multi_actions=acts, multi_logits=logits,
             multi_value=value, multi_logprob=lp, multi_entropy=ent, multi_nvec=NVEC_MULTI)
    tr.close()
    save("agent", **d)

    # ---- GAE
    d = {}
    for tag, T, N in (("a", 32, 8), ("b", 16, 33)):
        r = np.where(rng.uniform(size=(T, N)) < 0.05, -1.0, 1.0).astype(np.float32)
        v = rng.standard_normal((T, N)).astype(np.float32)
        dn = (rng.uniform(size=(T, N)) < 0.08).astype(np.float32)
        nv = rng.standard_normal(N).astype(np.float32)
        nd = (rng.uniform(size=N) < 0.2).astype(np.float32)
        # use_gae=false cannot be pinned: the reference's own branch (PPO_Discrete.cpp:309-329) throws
        # c10::Error at t=T-1 ("output with shape [N] doesn't match the broadcast shape [1, N]")
        # because next_value is [1,N] and is assigned into returns[t] ([N]).
        adv, ret = ref.gae(r, v, dn, nv, nd, np.float32(0.98), np.float32(0.95), use_gae=True)
        d.update({f"{tag}_adv_gae1": adv, f"{tag}_ret_gae1": ret})
        d.update({f"{tag}_rewards": r, f"{tag}_values": v, f"{tag}_dones": dn, f"{tag}_next_value": nv,
                  f"{tag}_next_done": nd})
    save("gae", gamma=np.float32(0.98), lam=np.float32(0.95), **d)

    # ---- minibatch update: 3 consecutive optimiser steps (loss, grads, clip, AdamW)
    def update_case(tag, O, K, nvec, masked, ent_coef, clip_vloss, norm_adv, seed):
        d = {}
        tr = ref.trainer(O, K, lr=1e-3, seed=seed)
        if nvec is not None:
            tr.set_nvec(nvec)
        heads = 1 if nvec is None else len(nvec)
        nv = np.array([K], np.int64) if nvec is None else nvec
        params = tr.get_params()
        # make the policy non-trivial so ratio/clip branches are exercised
        params[-(K * 64 + K):] *= 30.0
        tr.set_params(params)
        d["params0"] = params
        mb = 64
        for step in range(3):
            obs = rng.standard_normal((mb, O)).astype(np.float32)
            offs = np.concatenate([[0], np.cumsum(nv)[:-1]])
            mk = None
            if masked:
                mk = (rng.uniform(size=(mb, K)) < 0.7).astype(np.uint8)
                mk[:, offs] = 1
            acts = np.zeros((mb, heads), np.int64)
            for r in range(mb):
                for h, (o, k) in enumerate(zip(offs, nv)):
                    valid = np.arange(k) if mk is None else np.flatnonzero(mk[r, o:o + k])
                    acts[r, h] = rng.choice(valid)
            a_in = acts if masked else acts[:, 0]
            lp_now, _, val_now = tr.evaluate(obs, a_in, mask=mk)
            old_lp = (lp_now + rng.normal(0, 0.25, mb)).astype(np.float32)     # ratios spread around 1
            old_val = (val_now + rng.normal(0, 0.3, mb)).astype(np.float32)
            adv = rng.standard_normal(mb).astype(np.float32) * 2 + 0.3
            ret = (old_val + adv).astype(np.float32)
            lr = np.float32(1e-3 * (1 - step / 10))
            diag, grads = tr.minibatch_step(obs, a_in, old_lp, adv, ret, old_val, lr, mask=mk,
                                            clip_coef=np.float32(0.2), norm_adv=norm_adv, clip_vloss=clip_vloss,
                                            ent_coef=np.float32(ent_coef), vf_coef=np.float32(0.5),
                                            max_grad_norm=np.float32(0.5))
            d.update({f"s{step}_obs": obs, f"s{step}_actions": acts, f"s{step}_old_logprob": old_lp,
                      f"s{step}_old_val": old_val, f"s{step}_adv": adv, f"s{step}_ret": ret, f"s{step}_lr": lr,
                      f"s{step}_diag": diag, f"s{step}_grads": grads, f"s{step}_params_after": tr.get_params()})
            if mk is not None:
                d[f"s{step}_mask"] = mk
        tr.close()
        save("update_" + tag, O=O, K=K, nvec=nv, masked=int(masked), ent_coef=np.float32(ent_coef),
             clip_vloss=int(clip_vloss), norm_adv=int(norm_adv), **d)

    update_case("discrete", 4, 2, None, False, 0.01, True, True, 11)
    update_case("discrete_noclip", 4, 2, None, False, 0.0, False, False, 12)
    update_case("masked", 2, 3, np.array([3], np.int64), True, 0.01, True, True, 13)
    update_case("multihead", 2, 21, NVEC_MULTI, True, 0.01, True, True, 14)


if __name__ == "__main__":
    main()
