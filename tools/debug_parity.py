"""Stage-by-stage parity report of the CUDA hands step against the fp64 oracle (GPU box only).

Usage: python tools/debug_parity.py [scenario ...]   scenarios: free, curl, random, replay
For every step the GPU env is first synchronised to the oracle's (qpos, qvel, warmstart), so
each line is a SINGLE-SUBSTEP error; the last column is the free-running drift of a second
GPU env that is never re-synchronised.
"""
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))

import torch  # noqa: E402

from diffusion_piano_b200 import music, suite  # noqa: E402
from oracle.task_oracle import PianoWithShadowHandsOracle  # noqa: E402


def relerr(a, b):
    return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-12))


def run(name, ctrl_fn, steps, gravcomp=False, dt=0.005, verbose_every=1, max_contacts=0):
    task = suite.PianoWithShadowHands(music.load("TwinkleTwinkleRousseau"), trim_silence=True,
                                      n_steps_lookahead=1, control_timestep=dt, gravity_compensation=gravcomp)
    env = suite.BatchedEnvironment(task, 2, canonical_actions=False, auto_reset=False, max_contacts=max_contacts)
    ora = PianoWithShadowHandsOracle(task)
    env.reset()
    ora.reset()
    nv = 140
    dev = "cuda"
    qpos = torch.zeros((2, nv), device=dev); qvel = torch.zeros((2, nv), device=dev); warm = torch.zeros((2, nv), device=dev)
    qacc = torch.zeros((2, nv), device=dev); af = torch.zeros((2, 44), device=dev); av = torch.zeros((2, 44), device=dev)
    sx = torch.zeros((2, 10, 3), device=dev); cnt = torch.zeros((2, 8), dtype=torch.int32, device=dev)
    print(f"=== {name}: dt={dt} gravcomp={gravcomp}")
    print(" step | ncon g/o | nefc g/o | it g/o | qacc rel hand | qacc rel key | qpos abs | qvel abs | frc abs | site abs | rew g/o | drift qpos")
    worst = dict(qacc=0.0, qpos=0.0)
    for t in range(steps):
        a = ctrl_fn(t, ora)
        # sync env 0 to the oracle state; env 1 runs free
        env.engine.get_state(qpos=qpos, qvel=qvel, warm=warm)
        qpos[0] = torch.from_numpy(ora.physics.qpos.astype(np.float32)).to(dev)
        qvel[0] = torch.from_numpy(ora.physics.qvel.astype(np.float32)).to(dev)
        warm[0] = torch.from_numpy(ora.physics.qacc_warmstart.astype(np.float32)).to(dev)
        env.engine.set_state(qpos=qpos, qvel=qvel, warm=warm)
        torch.cuda.synchronize()
        acts = np.stack([a, a]).astype(np.float32)
        obs, rew, done = env.step(acts)
        obs_o, r_o, d_o = ora.step(a.astype(np.float32).astype(np.float64))
        env.engine.get_state(qpos=qpos, qvel=qvel, warm=warm)
        env.engine.get_diag(qacc=qacc, actuator_force=af, actuator_velocity=av, site_xpos=sx, counters=cnt)
        torch.cuda.synchronize()
        p = ora.physics
        qp, qv, qa = qpos.cpu().numpy(), qvel.cpu().numpy(), qacc.cpu().numpy()
        c = cnt.cpu().numpy()
        e_hand = relerr(qa[0, 88:], p.qacc[88:]); e_key = relerr(qa[0, :88], p.qacc[:88])
        e_qpos = float(np.abs(qp[0] - p.qpos).max()); e_qvel = float(np.abs(qv[0] - p.qvel).max())
        e_frc = float(np.abs(af.cpu().numpy()[0] - p.actuator_force).max())
        e_site = float(np.abs(sx.cpu().numpy()[0] - p.site_xpos).max())
        drift = float(np.abs(qp[1] - p.qpos).max())
        worst["qacc"] = max(worst["qacc"], relerr(qa[0], p.qacc)); worst["qpos"] = max(worst["qpos"], e_qpos)
        if t % verbose_every == 0 or not np.isfinite(e_qpos) or relerr(qa[0], p.qacc) > 1e-3:
            print(f"{t:5d} | {c[0,0]:3d}/{p.ncon:3d} | {c[0,1]:3d}/{p.nefc:3d} | {c[0,2]:2d}/{p.solver_iter:2d} | {e_hand:9.2e} | {e_key:9.2e} | "
                  f"{e_qpos:8.1e} | {e_qvel:8.1e} | {e_frc:7.1e} | {e_site:7.1e} | {rew[0]:.5f}/{r_o:.5f} | {drift:8.1e} | drop {c[0,3]} lr {c[0,7]} ls {c[0,5]}")
        if not np.isfinite(qp).all():
            print("non-finite state; worst hand dof:", np.argmax(np.abs(qa[0] - p.qacc)))
            break
    print(f"--- {name}: worst single-step qacc rel {worst['qacc']:.2e}, qpos abs {worst['qpos']:.2e}")
    env.close()
    return worst


def main():
    which = sys.argv[1:] or ["free", "curl", "random"]
    rng = np.random.RandomState(0)
    if "free" in which:
        run("free fall to zero ctrl", lambda t, o: np.zeros(45), 12)
        run("free, gravcomp", lambda t, o: np.zeros(45), 6, gravcomp=True)
    if "curl" in which or "curl_lr" in which:
        def curl(t, o):
            a = np.zeros(45)
            names = o.model.names["actuator"]
            a[names.index("rh_shadow_hand/rh_A_WRJ1")] = 0.45
            a[names.index("rh_shadow_hand/rh_A_FFJ3")] = 1.2
            a[22 + names.index("rh_shadow_hand/rh_A_WRJ1")] = 0.45
            a[22 + names.index("rh_shadow_hand/rh_A_MFJ3")] = 1.2
            return a
        if "curl" in which:
            run("curl fingers onto keys", curl, 160, verbose_every=8)
    if "curl_lr" in which:
        run("curl, all contacts low-rank (test hook)", curl, 160, verbose_every=4, max_contacts=256)
    for w in which:
      if w.startswith("random"):
        rng = np.random.RandomState(int(w[6:].rstrip("g") or 0))
        lo = hi = None

        def rnd(t, o):
            nonlocal lo, hi
            if lo is None:
                lo = np.concatenate([o.act_lo, [0.0]]); hi = np.concatenate([o.act_hi, [1.0]])
            if t % 10 == 0:
                rnd.a = lo + rng.uniform(0, 1, 45) * (hi - lo)
            return rnd.a
        run("random ctrl (held 10 substeps) " + w, rnd, 150, verbose_every=10, gravcomp=w.endswith("g"))


if __name__ == "__main__":
    main()
