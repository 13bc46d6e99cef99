"""Records the LAYOUT of a checkpoint written by the reference's own `ppo_v2.PPOAgent.save_checkpoint`
(ppo_v2.py:311-325) after one `update`: top-level keys, state_dict keys in order with shapes, optimiser
`param_groups` / per-parameter state keys, scheduler state keys -> tests/golden/ppo_checkpoint_layout.json.
Runs HERE (CPU) by importing /root/reference unmodified (wandb stubbed, ReduceLROnPlateau(verbose=) shimmed, as in
tools/make_ppo_golden.py); the GPU test compares the files `PPOAgent.save_checkpoint` of this repo writes with it.
usage: python tools/make_ppo_checkpoint_layout.py"""
import json
import sys
import types
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT / "tools"))
from make_ppo_golden import ppo_fixture_inputs  # noqa: E402


def main():
    import torch

    sys.modules.setdefault("wandb", types.ModuleType("wandb"))
    sys.path.insert(0, "/root/reference")
    real = torch.optim.lr_scheduler.ReduceLROnPlateau

    class Plateau(real):
        def __init__(self, *a, verbose=None, **k):
            super().__init__(*a, **k)

    torch.optim.lr_scheduler.ReduceLROnPlateau = Plateau
    import ppo_v2  # the reference, unmodified

    S, A, N = 24, 45, 64
    _, _, b = ppo_fixture_inputs(3, S, A, N)
    agent = ppo_v2.PPOAgent(S, A, batch_size=32, ppo_epochs=1, device="cpu", checkpoint_dir="/tmp/ppo_layout_ckpt",
                            use_wandb=False)
    agent.update(b["states"], b["actions"], b["rewards"].astype(np.float32), b["log_probs"], b["next_states"], b["dones"])
    agent.save_checkpoint(7, {"mean_rewards": [1.0]})
    ck = torch.load("/tmp/ppo_layout_ckpt/checkpoint_episode_7.pt", map_location="cpu", weights_only=False)

    def opt_layout(sd):
        g = sd["param_groups"][0]
        st = sd["state"][g["params"][0]]
        return {"n_groups": len(sd["param_groups"]), "group_keys": sorted(g), "params": list(g["params"]),
                "state_keys": sorted(st), "step": float(st["step"]), "step_is_tensor": torch.is_tensor(st["step"]),
                "betas": list(g["betas"]), "eps": g["eps"], "weight_decay": g["weight_decay"],
                "shapes": [list(sd["state"][i]["exp_avg"].shape) for i in g["params"]]}

    out = {"torch": torch.__version__, "keys": list(ck), "state_dim": S, "action_dim": A,
           "actor_state_dict": [[k, list(v.shape)] for k, v in ck["actor_state_dict"].items()],
           "critic_state_dict": [[k, list(v.shape)] for k, v in ck["critic_state_dict"].items()],
           "actor_parameters": [n for n, _ in agent.actor.named_parameters()],
           "critic_parameters": [n for n, _ in agent.critic.named_parameters()],
           "actor_optimizer": opt_layout(ck["actor_optimizer_state_dict"]),
           "critic_optimizer": opt_layout(ck["critic_optimizer_state_dict"]),
           "actor_scheduler_keys": sorted(ck["actor_scheduler_state_dict"]),
           "critic_scheduler_keys": sorted(ck["critic_scheduler_state_dict"]),
           "actor_scheduler": {k: ck["actor_scheduler_state_dict"][k] for k in ("mode", "factor", "patience", "best", "num_bad_epochs")},
           "critic_scheduler": {k: ck["critic_scheduler_state_dict"][k] for k in ("mode", "factor", "patience", "num_bad_epochs")}}
    p = ROOT / "tests" / "golden" / "ppo_checkpoint_layout.json"
    p.write_text(json.dumps(out, indent=1, default=float))
    print("wrote", p, p.stat().st_size, "bytes")


if __name__ == "__main__":
    main()
