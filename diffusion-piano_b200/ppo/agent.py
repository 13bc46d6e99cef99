"""Device-resident PPO v2 agent: same constructor, methods and update rule as the reference's
`PPOAgent` (`/root/reference/ppo_v2.py:134-334`), with everything kept in HBM.

What differs from the reference, by design:
* `select_actions` / `update` take and return CUDA tensors (numpy in -> numpy out is kept for drop-in use);
  nothing inside `update` reads a value back to the host - the reference's per-minibatch `.item()` / wandb
  calls, the Python loop over the batch index ("GAE", :236-246) and the DataLoader are gone.
* One minibatch update (gather, actor and critic forward, losses, backward, clip, Adam - the two networks on
  two streams) is ONE CUDA graph, replayed `ppo_epochs * ceil(N / batch_size)` times; cursor, epoch, RNG
  counter, Adam step counts, learning rates, the plateau schedulers and the RunningMeanStd live in a device
  control block (`ppo_ctrl`, include/ppo_kernels.h).
* fp32 arithmetic as the reference's default (`torch.backends.cuda.matmul.allow_tf32=False`): the minibatch GEMMs
  are this library's small-M FFMA kernels (cuBLAS needs 4.7 - 8 us for a 128-row SGEMM inside the graph), whole-batch
  inference uses cuBLAS SGEMM, every other operation is a hand-written kernel of libppo_kernels.so. No autograd.
* Shuffling is a keyed Feistel permutation computed in the gather kernel, dropout masks are Philox draws: the
  same distributions as the reference's, not the same streams (tests inject explicit permutations / masks).
* Several GPUs (`process_group`): every rank steps its own env shard; the reward / value vectors are
  all-gathered so that the normaliser, the index-wise recursion and the advantage statistics see the global
  batch exactly as a single process would; each rank then runs minibatches of `batch_size` rows of ITS shard and
  the gradients are averaged with one NCCL all-reduce per network per minibatch (captured in the graph).
"""

from __future__ import annotations

import ctypes
from pathlib import Path
from typing import Dict, List, Optional, Sequence

import numpy as np
import torch

from . import _kernels, dist

HIDDEN = 256
ACTOR_LINEARS, ACTOR_NORMS = (0, 3, 6, 9), (2, 5, 8)
CRITIC_LINEARS, CRITIC_NORMS = (0, 4, 8, 11), (2, 6, 10)
DROPOUT_P = 0.1


def _reference_init(state_dim: int, action_dim: int, seed: Optional[int]):
    """Initial parameters drawn exactly as the reference's modules draw them (ppo_v2.py:51-105): default
    nn.Linear / nn.LayerNorm init for the actor, orthogonal(gain 0.01) + zero bias for the critic's Linears."""
    nn = torch.nn
    gen_state = torch.random.get_rng_state()
    if seed is not None:
        torch.manual_seed(seed)
    try:
        h = HIDDEN
        actor = nn.Sequential(nn.Linear(state_dim, h), nn.ReLU(), nn.LayerNorm(h), nn.Linear(h, h), nn.ReLU(),
                              nn.LayerNorm(h), nn.Linear(h, h), nn.ReLU(), nn.LayerNorm(h), nn.Linear(h, action_dim),
                              nn.Tanh())
        critic = nn.Sequential(nn.Linear(state_dim, h), nn.ReLU(), nn.LayerNorm(h), nn.Dropout(0.1), nn.Linear(h, h),
                               nn.ReLU(), nn.LayerNorm(h), nn.Dropout(0.1), nn.Linear(h, h // 2), nn.ReLU(),
                               nn.LayerNorm(h // 2), nn.Linear(h // 2, 1))
        for m in critic.modules():
            if isinstance(m, nn.Linear):
                nn.init.orthogonal_(m.weight, gain=0.01)
                nn.init.constant_(m.bias, 0)
    finally:
        if seed is not None:
            torch.random.set_rng_state(gen_state)
    a = {f"network.{k}": v.detach().numpy().copy() for k, v in actor.state_dict().items()}
    a["log_std"] = np.full(action_dim, -1.0, np.float32)
    c = {f"network.{k}": v.detach().numpy().copy() for k, v in critic.state_dict().items()}
    return a, c


class _Net:
    """One network's flat parameter / gradient / Adam buffers with per-tensor views, and its activations."""

    def __init__(self, dims: Sequence[int], linears, norms, dropout_after, extra: Dict[str, int], device):
        self.dims, self.linears, self.norms, self.dropout_after = list(dims), linears, norms, dropout_after
        self.names: List[str] = []
        self.shapes: Dict[str, tuple] = {}
        for blk in range(4):
            li = linears[blk]
            self.shapes[f"network.{li}.weight"] = (dims[blk + 1], dims[blk])
            self.shapes[f"network.{li}.bias"] = (dims[blk + 1],)
            if blk < 3:
                ni = norms[blk]
                self.shapes[f"network.{ni}.weight"] = (dims[blk + 1],)
                self.shapes[f"network.{ni}.bias"] = (dims[blk + 1],)
        for k, n in extra.items():
            self.shapes[k] = (n,)
        self.offsets, off = {}, 0
        for k, shp in self.shapes.items():
            self.offsets[k] = off
            off += (int(np.prod(shp)) + 3) // 4 * 4  # 16-byte aligned tensors
        self.size = off
        z = lambda: torch.zeros(self.size, dtype=torch.float32, device=device)  # noqa: E731
        self.theta, self.grad, self.m, self.v = z(), z(), z(), z()
        self.p = {k: self._view(self.theta, k) for k in self.shapes}
        self.g = {k: self._view(self.grad, k) for k in self.shapes}
        self.device = device
        self.scratch = torch.zeros(3 * HIDDEN * 256 + 1024, dtype=torch.float32, device=device)
        self.ticket = torch.zeros(4, dtype=torch.int32, device=device)
        self.wstream = torch.cuda.Stream(device=device)  # weight-gradient GEMMs
        self.cstream = torch.cuda.Stream(device=device)  # bias / LayerNorm parameter gradients (column sums)
        self.train: Dict[int, dict] = {}

    def _view(self, flat, k):
        n = int(np.prod(self.shapes[k]))
        return flat[self.offsets[k]:self.offsets[k] + n].view(self.shapes[k])

    def load(self, params: Dict[str, np.ndarray]):
        for k in self.shapes:
            self.p[k].copy_(torch.as_tensor(np.asarray(params[k], np.float32)).view(self.shapes[k]))

    def state_dict(self) -> Dict[str, torch.Tensor]:
        """Keys in the reference modules' `state_dict()` order: a module's own parameters (`log_std`) first."""
        keys = [k for k in self.p if not k.startswith("network.")] + [k for k in self.p if k.startswith("network.")]
        return {k: self.p[k].detach().cpu().clone() for k in keys}

    def buffers(self, rows: int, training: bool) -> dict:
        """Activation buffers for `rows` rows (allocated once per distinct row count)."""
        key = rows if training else -rows
        if key not in self.train:
            f = lambda *s: torch.empty(*s, dtype=torch.float32, device=self.device)  # noqa: E731
            b = {"y": [f(rows, self.dims[i + 1]) for i in range(4)], "a": [f(rows, self.dims[i + 1]) for i in range(3)]}
            if training:
                b["stats"] = [f(rows, 2) for _ in range(3)]
                b["h"] = [f(rows, self.dims[i + 1]) for i in range(3)]  # post-ReLU values (LayerNorm backward)
                b["mask"] = [torch.empty(rows, self.dims[i + 1], dtype=torch.uint8, device=self.device)
                             if self.dropout_after[i] else None for i in range(3)]
                b["da"] = [f(rows, self.dims[i + 1]) for i in range(3)]
                b["dy"] = [f(rows, self.dims[i + 1]) for i in range(4)]
            self.train[key] = b
        return self.train[key]


class PPOAgent:
    def __init__(self, state_dim, action_dim, lr=1e-4, gamma=0.99, epsilon=0.2, entropy_coef=0.01, value_coef=1.0,
                 max_grad_norm=0.5, ppo_epochs=10, batch_size=64, device="cuda", checkpoint_dir="checkpoints",
                 use_wandb=False, seed: Optional[int] = 0, process_group=None, dropout_p: float = DROPOUT_P,
                 use_graphs: bool = True, graph_scope: str = "update", gemm: str = "custom",
                 fuse_value_head: bool = True):
        if use_wandb:
            raise NotImplementedError("wandb logging is not part of the device path; read agent.stats() instead")
        if not torch.cuda.is_available():
            raise _kernels.PPOKernelError("the PPO update runs on a CUDA device only (no CPU fallback)")
        self.lib = _kernels.load()
        self.kernel_launches = 0  # launches of this library's kernels issued or replayed (bench accounting)
        self.device = torch.device(device if str(device) != "cuda" else f"cuda:{torch.cuda.current_device()}")
        self.state_dim, self.action_dim = int(state_dim), int(action_dim)
        self.gamma, self.epsilon, self.entropy_coef, self.value_coef = gamma, epsilon, entropy_coef, value_coef
        self.max_grad_norm, self.ppo_epochs, self.batch_size = max_grad_norm, ppo_epochs, batch_size
        self.lam, self.dropout_p, self.use_graphs = 0.95, float(dropout_p), use_graphs
        if graph_scope not in ("update", "minibatch"):
            raise ValueError("graph_scope must be 'update' or 'minibatch'")
        self.graph_scope = graph_scope
        if gemm not in ("custom", "cublas"):
            raise ValueError("gemm must be 'custom' or 'cublas'")
        self.gemm, self.small_m = gemm, 512
        # the value head with the critic's last Linear and its dA inside (ppo_value_head_fused): 25.0 ms per update
        # against 25.3 ms (B200, B = 4096). The other fusions that were built and measured slower (policy head with
        # its Linear inside, LayerNorm as a GEMM prologue, a cluster-wide Linear + LayerNorm kernel, a peer-memory
        # all-reduce) live under tools/experiments, not in the product library.
        self.fuse_value_head = bool(fuse_value_head)
        self.pg = process_group
        self.world = torch.distributed.get_world_size(process_group) if process_group is not None else 1
        self.rank = torch.distributed.get_rank(process_group) if process_group is not None else 0
        # the two networks' gradient all-reduces sit on concurrent branches of the minibatch graph: each gets its
        # own communicator so that neither can be ordered behind the other differently on different ranks
        self.pg_nets = [process_group, process_group]
        if self.world > 1:
            ranks = torch.distributed.get_process_group_ranks(process_group)
            self.pg_nets = [torch.distributed.new_group(ranks), torch.distributed.new_group(ranks)]
        h = HIDDEN
        self.actor = _Net((state_dim, h, h, h, action_dim), ACTOR_LINEARS, ACTOR_NORMS, (False, False, False),
                          {"log_std": action_dim}, self.device)
        self.critic = _Net((state_dim, h, h, h // 2, 1), CRITIC_LINEARS, CRITIC_NORMS, (True, True, False), {},
                           self.device)
        a, c = _reference_init(state_dim, action_dim, seed)
        self.actor.load(a)
        self.critic.load(c)
        if self.world > 1:  # every rank starts from rank 0's parameters
            torch.distributed.broadcast(self.actor.theta, 0, group=self.pg)
            torch.distributed.broadcast(self.critic.theta, 0, group=self.pg)
        self.ctrl = torch.zeros(ctypes.sizeof(_kernels.PPOCtrl), dtype=torch.uint8, device=self.device)
        # ranks share the seed for the schedule state but draw different noise / dropout / permutations
        self._k(self.lib.ppo_ctrl_init, self.ctrl, (seed or 0) * 1000003 + self.rank, lr, lr * 2)
        self.checkpoint_dir = Path(checkpoint_dir)
        self._side = torch.cuda.Stream(device=self.device)
        self._pref = torch.cuda.Stream(device=self.device)  # gather of the next minibatch
        self._graphs: Dict[tuple, torch.cuda.CUDAGraph] = {}
        self._batch: Dict[int, dict] = {}
        self._launches_per_graph: Dict[tuple, int] = {}

    # ------------------------------------------------------------------ helpers
    def _k(self, fn, *args):
        """Calls a C-ABI entry point on the current stream; tensors become device pointers."""
        conv = [a.data_ptr() if isinstance(a, torch.Tensor) else a for a in args]
        _kernels.check(fn(*conv, torch.cuda.current_stream(self.device).cuda_stream))
        self.kernel_launches += 1

    # GEMMs: this library's small-M kernels for minibatches, cuBLAS (torch.mm) for whole-batch inference
    def _mm_nt(self, a, b, out):  # out[M,N] = a[M,K] . b[N,K]^T
        if self.gemm == "custom" and a.shape[0] <= self.small_m:
            self._k(self.lib.ppo_gemm_nt, a, b, out, a.shape[0], b.shape[0], a.shape[1])
        else:
            torch.mm(a, b.t(), out=out)

    def _mm_nn(self, a, b, out):  # out[M,K] = a[M,N] . b[N,K]
        if self.gemm == "custom" and a.shape[0] <= self.small_m:
            self._k(self.lib.ppo_gemm_nn, a, b, out, a.shape[0], a.shape[1], b.shape[1])
        else:
            torch.mm(a, b, out=out)

    def _mm_tn(self, a, b, out):  # out[N,K] = a[M,N]^T . b[M,K]
        if self.gemm == "custom" and a.shape[0] <= self.small_m:
            self._k(self.lib.ppo_gemm_tn, a, b, out, a.shape[0], a.shape[1], b.shape[1])
        else:
            torch.mm(a.t(), b, out=out)

    def _block_fwd(self, net: _Net, blk: int, x, buf, training: bool, p: float, mask_in, rng_stream: int):
        li, ni = net.linears[blk], net.norms[blk]
        use_p = p if net.dropout_after[blk] else 0.0
        self._mm_nt(x, net.p[f"network.{li}.weight"], buf["y"][blk])
        self._k(self.lib.ppo_relu_ln_fwd, buf["y"][blk], net.p[f"network.{li}.bias"], net.p[f"network.{ni}.weight"],
                net.p[f"network.{ni}.bias"], buf["a"][blk], buf["h"][blk] if training else None,
                buf["stats"][blk] if training else None, x.shape[0], net.dims[blk + 1], use_p,
                mask_in if use_p > 0 else None,
                buf["mask"][blk] if (training and use_p > 0) else None, self.ctrl, rng_stream + blk)
        return buf["a"][blk]

    def _forward(self, net: _Net, x, training: bool, p: float = 0.0, masks=None, rng_stream: int = 0,
                 last: bool = True):
        """Three blocks + the last Linear (without its bias: the head kernels add it). Returns y4 (last=False: the
        fused head kernels compute the last Linear themselves)."""
        buf = net.buffers(x.shape[0], training)
        a, mi = x, 0
        for blk in range(3):
            m = None
            if masks is not None and net.dropout_after[blk]:
                m, mi = masks[mi], mi + 1
            a = self._block_fwd(net, blk, a, buf, training, p, m, rng_stream)
        if last:
            self._mm_nt(a, net.p[f"network.{net.linears[3]}.weight"], buf["y"][3])
        return buf["y"][3], buf

    def _backward(self, net: _Net, x, buf, dz, p: float, da_ready: bool = False):
        """dz = gradient with respect to the last Linear's output; fills net.grad (except the head's bias).
        da_ready: the fused head already wrote dA of the third block into buf["da"][2]."""
        chain = torch.cuda.current_stream(self.device)
        wstream = net.wstream  # weight-gradient GEMMs leave the dependent chain dA -> LN backward -> dA ...
        li = net.linears[3]
        wstream.wait_stream(chain)
        with torch.cuda.stream(wstream):
            self._mm_tn(dz, buf["a"][2], net.g[f"network.{li}.weight"])
        if not da_ready:
            self._mm_nn(dz, net.p[f"network.{li}.weight"], buf["da"][2])
        for blk in (2, 1, 0):
            li, ni = net.linears[blk], net.norms[blk]
            use_p = p if net.dropout_after[blk] else 0.0
            mask = buf["mask"][blk] if use_p > 0 else None
            fused = blk > 0 and self.gemm == "custom" and x.shape[0] <= self.small_m
            if fused:  # dy of this block and dA for the block below in one launch
                self._k(self.lib.ppo_ln_bwd_gemm_nn, buf["da"][blk], buf["h"][blk], buf["stats"][blk],
                        net.p[f"network.{ni}.weight"], mask, use_p, buf["dy"][blk], net.p[f"network.{li}.weight"],
                        buf["da"][blk - 1], x.shape[0], net.dims[blk + 1], net.dims[blk])
            else:
                self._k(self.lib.ppo_relu_ln_bwd_dy, buf["da"][blk], buf["h"][blk], buf["stats"][blk],
                        net.p[f"network.{ni}.weight"], mask, use_p, buf["dy"][blk], x.shape[0], net.dims[blk + 1])
            a_prev = x if blk == 0 else buf["a"][blk - 1]
            # parameter gradients of the block leave the chain: column sums and the weight-gradient GEMM
            wstream.wait_stream(chain)
            net.cstream.wait_stream(chain)
            with torch.cuda.stream(net.cstream):
                self._k(self.lib.ppo_relu_ln_colsums, buf["da"][blk], buf["h"][blk], buf["stats"][blk], mask, use_p,
                        buf["dy"][blk], net.g[f"network.{li}.bias"], net.g[f"network.{ni}.weight"],
                        net.g[f"network.{ni}.bias"], x.shape[0], net.dims[blk + 1])
            with torch.cuda.stream(wstream):
                self._mm_tn(buf["dy"][blk], a_prev, net.g[f"network.{li}.weight"])
            if blk > 0 and not fused:
                self._mm_nn(buf["dy"][blk], net.p[f"network.{li}.weight"], buf["da"][blk - 1])
        chain.wait_stream(wstream)
        chain.wait_stream(net.cstream)

    def _optim_step(self, net: _Net, which: int):
        if self.world > 1:
            torch.distributed.all_reduce(net.grad, group=self.pg_nets[which])
        self._k(self.lib.ppo_grad_sumsq, net.grad, net.size, net.scratch, net.ticket[1:], self.ctrl, which)
        self._k(self.lib.ppo_clip_adam, net.theta, net.grad, net.m, net.v, net.size, self.ctrl, which,
                self.max_grad_norm, 1.0 / self.world, 0.9, 0.999, 1e-5)

    def _batch_buffers(self, n: int) -> dict:
        if n not in self._batch:
            f = lambda *s: torch.empty(*s, dtype=torch.float32, device=self.device)  # noqa: E731
            ng = n * self.world
            self._batch[n] = dict(
                states=f(n, self.state_dim), next_states=f(n, self.state_dim), actions=f(n, self.action_dim),
                logp=f(n), packed=f(4, n), gathered=f(self.world, 4, n), glob=f(4, ng), rewards_n=f(ng),
                returns=f(ng), adv_raw=f(ng), adv=f(ng), values=f(n), next_values=f(n))
        return self._batch[n]

    def _minibatch_buffers(self, B: int, parity: int = 0) -> dict:
        key = ("mb", B, parity)
        if key not in self._batch:
            f = lambda *s: torch.empty(*s, dtype=torch.float32, device=self.device)  # noqa: E731
            self._batch[key] = dict(states=f(B, self.state_dim), actions=f(B, self.action_dim), logp=f(B), adv=f(B),
                                    ret=f(B), dz=f(B, self.action_dim), dv=f(B, 1), new_logp=f(B))
        return self._batch[key]

    # ------------------------------------------------------------------ one minibatch (graph body)
    def _gather(self, bb: dict, n: int, B: int, adv, returns, perm, perm_epochs: int, parity: int) -> dict:
        mb = self._minibatch_buffers(B, parity)
        self._k(self.lib.ppo_gather_minibatch, bb["states"], bb["actions"], bb["logp"], adv, returns, perm,
                perm_epochs, self.ctrl, 0, n, self.state_dim, self.action_dim, B, mb["states"], mb["actions"],
                mb["logp"], mb["adv"], mb["ret"])
        return mb

    def _minibatch(self, bb: dict, n: int, B: int, adv, returns, perm=None, perm_epochs: int = 0, masks=None,
                   mb: Optional[dict] = None, next_B: Optional[int] = None, parity: int = 0):
        """One minibatch update. `mb`: rows already gathered by the previous call's prefetch; `next_B`: gather the
        next minibatch (into the other buffer set) as soon as the cursor has moved, off both networks' chains.
        Returns the prefetched buffers (or None)."""
        main = torch.cuda.current_stream(self.device)
        if mb is None:
            mb = self._gather(bb, n, B, adv, returns, perm, perm_epochs, parity)
        else:
            main.wait_stream(self._pref)
        side = self._side
        side.wait_stream(main)
        next_mb = None
        # critic on the side stream (ppo_v2.py:274-282: critic first - the two steps are independent)
        with torch.cuda.stream(side):
            net = self.critic
            small = self.gemm == "custom" and B <= self.small_m
            fused = small and self.fuse_value_head
            y4, buf = self._forward(net, mb["states"], True, self.dropout_p, masks, rng_stream=1, last=not fused)
            # cursor / epoch / RNG counter were last read by the gather and the dropout above: bump them here,
            # off the actor's (longer) chain
            self._k(self.lib.ppo_advance, self.ctrl, n, B)
            if next_B is not None:
                self._pref.wait_stream(side)
                with torch.cuda.stream(self._pref):
                    next_mb = self._gather(bb, n, next_B, adv, returns, perm, perm_epochs, parity ^ 1)
            if fused:
                self._k(self.lib.ppo_value_head_fused, buf["a"][2], net.p["network.11.weight"],
                        net.p["network.11.bias"], mb["ret"], mb["dv"], buf["da"][2], net.g["network.11.bias"],
                        self.ctrl, B, net.dims[3])
            else:
                self._k(self.lib.ppo_value_head, y4, net.p["network.11.bias"], mb["ret"], mb["dv"],
                        net.g["network.11.bias"], self.ctrl, B)
            self._backward(net, mb["states"], buf, mb["dv"], self.dropout_p, da_ready=fused)
            self._optim_step(net, 1)
        net = self.actor
        y4, buf = self._forward(net, mb["states"], True)
        self._k(self.lib.ppo_policy_head, y4, net.p["network.9.bias"], net.p["log_std"], mb["actions"], mb["logp"],
                mb["adv"], self.epsilon, self.entropy_coef, mb["dz"], net.g["network.9.bias"], net.g["log_std"],
                mb["new_logp"], self.ctrl, net.scratch, net.ticket[2:], B, self.action_dim)
        self._backward(net, mb["states"], buf, mb["dz"], 0.0)
        self._optim_step(net, 0)
        main.wait_stream(side)
        return next_mb

    def _replay(self, key, sizes, bb, n, adv, returns):
        """Replays the CUDA graph of the minibatches `sizes` (one minibatch, or all of an update), capturing it on
        first use."""
        if key not in self._graphs:
            # warm up once eagerly on a side stream (cuBLAS workspaces, lazy module loads), with the control
            # block and the parameters restored afterwards so the warm-up is not a training step
            state = (self.ctrl, self.actor.theta, self.actor.m, self.actor.v, self.critic.theta, self.critic.m,
                     self.critic.v)
            saved = [t.clone() for t in state]
            s = torch.cuda.Stream(device=self.device)
            s.wait_stream(torch.cuda.current_stream(self.device))
            with torch.cuda.stream(s):
                for B in sorted(set(sizes)):
                    self._minibatch_buffers(B, 1)  # both buffer sets exist before the capture
                    self._minibatch(bb, n, B, adv, returns)
            torch.cuda.current_stream(self.device).wait_stream(s)
            for t, v in zip(state, saved):
                t.copy_(v)
            before = self.kernel_launches
            g = torch.cuda.CUDAGraph()
            # thread_local: the NCCL watchdog thread may query events while this thread captures
            with torch.cuda.graph(g, capture_error_mode="thread_local" if self.world > 1 else "global"):
                mb = None
                for i, B in enumerate(sizes):
                    mb = self._minibatch(bb, n, B, adv, returns, mb=mb,
                                         next_B=sizes[i + 1] if i + 1 < len(sizes) else None, parity=i & 1)
            self._launches_per_graph[key] = self.kernel_launches - before
            self.kernel_launches = before
            self._graphs[key] = g  # capture does not execute: nothing to restore
        self._graphs[key].replay()
        self.kernel_launches += self._launches_per_graph[key]

    # ------------------------------------------------------------------ public API (ppo_v2.py:202-334)
    def select_actions(self, states, noise=None):
        """Batch of states -> (actions, log_probs); CUDA tensors in -> CUDA tensors out (reused buffers),
        numpy in -> numpy out as the reference."""
        as_numpy = isinstance(states, np.ndarray)
        x = torch.as_tensor(states, dtype=torch.float32, device=self.device) if as_numpy else states
        n = x.shape[0]
        bb = self._batch_buffers(n)
        y4, _ = self._forward(self.actor, x, False)
        self._k(self.lib.ppo_sample_actions, y4, self.actor.p["network.9.bias"], self.actor.p["log_std"], noise,
                bb["actions"], bb["logp"], self.ctrl, n, self.action_dim)
        self._k(self.lib.ppo_advance, self.ctrl, 0, 0)
        if as_numpy:
            return bb["actions"].cpu().numpy(), bb["logp"].cpu().numpy()
        return bb["actions"], bb["logp"]

    def update(self, states, actions, rewards, log_probs, next_states, dones, perms=None, value_masks=None,
               next_value_masks=None, minibatch_masks=None):
        """One call per control step, as `parallelized_base_v2.py:147-154`. Explicit `perms` [E,N] int32 and
        dropout keep-masks (uint8 tensors) are test hooks and run the minibatches eagerly."""
        dev = self.device
        t = lambda v, dt=torch.float32: torch.as_tensor(np.asarray(v) if not isinstance(v, torch.Tensor) else v,  # noqa: E731
                                                        device=dev).to(dt)
        n = len(rewards)
        bb = self._batch_buffers(n)
        bb["states"].copy_(t(states))
        bb["next_states"].copy_(t(next_states))
        if not (isinstance(actions, torch.Tensor) and actions.data_ptr() == bb["actions"].data_ptr()):
            bb["actions"].copy_(t(actions))
        if not (isinstance(log_probs, torch.Tensor) and log_probs.data_ptr() == bb["logp"].data_ptr()):
            bb["logp"].copy_(t(log_probs))
        packed = bb["packed"]
        packed[0].copy_(t(rewards))
        packed[1].copy_(t(dones))
        explicit = perms is not None or value_masks is not None or minibatch_masks is not None
        p = self.dropout_p
        # critic on the whole batch, in training mode (the reference never calls .eval(): dropout is active)
        for row, x, masks, rs in ((2, bb["states"], value_masks, 11), (3, bb["next_states"], next_value_masks, 21)):
            y4, _ = self._forward(self.critic, x, False, p, masks, rng_stream=rs)
            self._k(self.lib.ppo_add_bias, y4, self.critic.p["network.11.bias"], packed[row], n)
        if self.world > 1:
            glob = dist.gather_global_batch(packed, bb["gathered"], bb["glob"], self.pg)
        else:
            glob = packed
        ng = n * self.world
        self._k(self.lib.ppo_prepare_batch, glob[0], glob[1], glob[2], glob[3], self.gamma, self.lam, bb["rewards_n"],
                bb["returns"], bb["adv_raw"], bb["adv"], self.ctrl, ng)
        mine = dist.shard_slice(self.rank, n)
        adv, returns = bb["adv"][mine], bb["returns"][mine]
        self._k(self.lib.ppo_begin_update, self.ctrl)
        perm_t = None if perms is None else torch.as_tensor(np.asarray(perms), device=dev).to(torch.int32).contiguous()
        sizes = [min(self.batch_size, n - start) for _ in range(self.ppo_epochs)
                 for start in range(0, n, self.batch_size)]
        if explicit or not self.use_graphs:
            mb = None
            for mb_index, B in enumerate(sizes):
                masks = None if minibatch_masks is None else minibatch_masks[mb_index]
                mb = self._minibatch(bb, n, B, adv, returns, perm_t, 0 if perm_t is None else perm_t.shape[0], masks,
                                     mb=mb, next_B=sizes[mb_index + 1] if mb_index + 1 < len(sizes) else None,
                                     parity=mb_index & 1)
        elif self.graph_scope == "update":  # every minibatch of the update in one graph launch
            self._replay((n, "update", adv.data_ptr(), returns.data_ptr()), sizes, bb, n, adv, returns)
        else:
            for B in sizes:
                self._replay((n, B, adv.data_ptr(), returns.data_ptr()), [B], bb, n, adv, returns)
        if self.world > 1:  # one critic-loss value for all ranks: the schedulers must step identically everywhere
            dist.sync_scheduler_metric(self.ctrl, _kernels.PPOCtrl.critic_loss.offset, self.world, self.pg)
        self._k(self.lib.ppo_plateau_step, self.ctrl, 0.5, 100)
        return bb

    def close(self):
        """Drops the captured graphs (they hold NCCL kernels: release them before destroying the process group)."""
        torch.cuda.synchronize(self.device)
        self._graphs.clear()

    # ------------------------------------------------------------------ introspection / checkpoints
    def read_ctrl(self) -> _kernels.PPOCtrl:
        raw = self.ctrl.cpu().numpy().tobytes()
        return _kernels.PPOCtrl.from_buffer_copy(raw)

    def write_ctrl(self, c: "_kernels.PPOCtrl") -> None:
        torch.cuda.synchronize(self.device)
        self.ctrl.copy_(torch.frombuffer(bytearray(bytes(c)), dtype=torch.uint8))

    def stats(self) -> Dict[str, float]:
        c = self.read_ctrl()
        return dict(actor_loss=c.actor_loss, critic_loss=c.critic_loss, entropy=c.entropy, mean_ratio=c.mean_ratio,
                    value_predictions=c.mean_value, mean_reward_normalized=c.mean_reward_n, lr_actor=c.lr[0],
                    lr_critic=c.lr[1], adam_steps=c.adam_t[0], grad_norm_actor=float(np.sqrt(c.norm2[0])) / self.world,
                    grad_norm_critic=float(np.sqrt(c.norm2[1])) / self.world, reward_mean=c.rms_mean,
                    reward_var=c.rms_var)

    def actor_state_dict(self):
        return self.actor.state_dict()

    def critic_state_dict(self):
        return self.critic.state_dict()

    # ---- checkpoints in the reference's layout (ppo_v2.py:311-336)
    _SCHED = (("actor", 0, "max"), ("critic", 1, "min"))

    @staticmethod
    def _param_order(net: "_Net") -> List[str]:
        """`module.parameters()` order of the reference's Actor / Critic: a module's own parameters (`log_std`)
        come before its children's (`network.*`, in Sequential order)."""
        own = [k for k in net.shapes if not k.startswith("network.")]
        return own + [k for k in net.shapes if k.startswith("network.")]

    def _torch_optimizer(self, net: "_Net", which: int, c: "_kernels.PPOCtrl"):
        """A torch Adam + ReduceLROnPlateau pair carrying this network's state: their `state_dict()`s are what the
        reference's `load_checkpoint` feeds to `load_state_dict` (same torch, same keys)."""
        order = self._param_order(net)
        params = [torch.nn.Parameter(net.p[k].detach().cpu().clone()) for k in order]
        opt = torch.optim.Adam(params, lr=float(c.lr[which]), betas=(0.9, 0.999), eps=1e-5)
        if c.adam_t[which] > 0:
            for prm, k in zip(params, order):
                opt.state[prm] = {"step": torch.tensor(float(c.adam_t[which])),
                                  "exp_avg": net._view(net.m, k).detach().cpu().clone(),
                                  "exp_avg_sq": net._view(net.v, k).detach().cpu().clone()}
        mode = self._SCHED[which][2]
        sched = torch.optim.lr_scheduler.ReduceLROnPlateau(opt, mode=mode, factor=0.5, patience=100)
        sched.best = float(c.best[which])
        sched.num_bad_epochs = int(c.num_bad[which])
        sched._last_lr = [float(c.lr[which])]
        return opt, sched

    def save_checkpoint(self, episode, rewards):
        """The reference's checkpoint (ppo_v2.py:311-325), key for key: network state_dicts, torch Adam state_dicts
        (per-parameter `step` / `exp_avg` / `exp_avg_sq`, `param_groups` with the current lr) and
        ReduceLROnPlateau state_dicts (`best`, `num_bad_epochs`), so that `ppo_v2.PPOAgent.load_checkpoint` reads
        it. One extra key, `b200_state`, holds what the reference does not save and an exact resume here needs
        (reward normaliser, RNG counter, epoch cursor); the rank-specific seed is NOT part of it."""
        self.checkpoint_dir.mkdir(parents=True, exist_ok=True)
        c = self.read_ctrl()
        ckpt = {"episode": episode, "actor_state_dict": self.actor_state_dict(),
                "critic_state_dict": self.critic_state_dict()}
        pairs = [(name, self._torch_optimizer(net, which, c))
                 for (name, which, _), net in zip(self._SCHED, (self.actor, self.critic))]
        for name, (opt, _) in pairs:  # the reference's key order: both optimisers, then both schedulers
            ckpt[f"{name}_optimizer_state_dict"] = opt.state_dict()
        for name, (_, sched) in pairs:
            ckpt[f"{name}_scheduler_state_dict"] = sched.state_dict()
        ckpt["rewards"] = rewards
        ckpt["b200_state"] = {"rms_mean": c.rms_mean, "rms_var": c.rms_var, "rms_count": c.rms_count,
                              "rng_counter": int(c.rng_counter), "epoch": int(c.epoch)}
        path = self.checkpoint_dir / f"checkpoint_episode_{episode}.pt"
        torch.save(ckpt, path)
        return path

    def save_model(self, path):
        torch.save({"actor_state_dict": self.actor_state_dict(), "critic_state_dict": self.critic_state_dict()}, path)

    def load_checkpoint(self, path):
        """Reads a checkpoint written by `save_checkpoint` here OR by the reference's `ppo_v2.PPOAgent` (same
        keys): networks, Adam moments and step counts, learning rates, plateau-scheduler state. Optimiser state
        that is present but cannot be mapped onto the flat buffers raises instead of being dropped. The control
        block's seed stays this rank's own (ranks must not share noise / dropout / permutation streams)."""
        ckpt = torch.load(path, map_location="cpu", weights_only=False)
        self.actor.load({k: v.numpy() for k, v in ckpt["actor_state_dict"].items()})
        self.critic.load({k: v.numpy() for k, v in ckpt["critic_state_dict"].items()})
        c = self.read_ctrl()
        for (name, which, _), net in zip(self._SCHED, (self.actor, self.critic)):
            osd = ckpt.get(f"{name}_optimizer_state_dict")
            if osd is not None:
                order = self._param_order(net)
                ids = osd["param_groups"][0]["params"]
                if len(osd["param_groups"]) != 1 or len(ids) != len(order):
                    raise ValueError(f"{name} optimiser state: {len(ids)} parameters in {len(osd['param_groups'])} "
                                     f"groups, expected {len(order)} in one group")
                c.lr[which] = float(osd["param_groups"][0]["lr"])
                net.m.zero_(); net.v.zero_()
                steps = set()
                for pid, k in zip(ids, order):
                    st = osd["state"].get(pid)
                    if st is None:
                        continue
                    if tuple(st["exp_avg"].shape) != tuple(net.shapes[k]):
                        raise ValueError(f"{name} optimiser state of {k}: shape {tuple(st['exp_avg'].shape)}, "
                                         f"expected {net.shapes[k]}")
                    net._view(net.m, k).copy_(st["exp_avg"])
                    net._view(net.v, k).copy_(st["exp_avg_sq"])
                    steps.add(int(float(st["step"])))
                if len(steps) > 1:
                    raise ValueError(f"{name} optimiser state: parameters at different Adam steps {sorted(steps)}")
                c.adam_t[which] = steps.pop() if steps else 0
            ssd = ckpt.get(f"{name}_scheduler_state_dict")
            if ssd is not None:
                c.best[which] = float(ssd["best"])
                c.num_bad[which] = int(ssd["num_bad_epochs"])
        extra = ckpt.get("b200_state")
        if extra is not None:
            c.rms_mean, c.rms_var, c.rms_count = extra["rms_mean"], extra["rms_var"], extra["rms_count"]
            c.rng_counter, c.epoch = extra["rng_counter"], extra["epoch"]
        self.write_ctrl(c)
        return ckpt.get("episode", 0), ckpt.get("rewards")
