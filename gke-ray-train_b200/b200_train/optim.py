"""b200_train.AdamW — drop-in for torch.optim.AdamW on the reference's step tail.

ref: ray-jobs/pytorch_llm_ray.py:236 `optim.AdamW(model.parameters(), lr=lr, weight_decay=0.01)`, :273
`optimizer.zero_grad(set_to_none=True)`, :278 `optimizer.step()`, :302 `optimizer.state_dict()`; the HF Trainer
AdamW with two param_groups configured by ref: ray-jobs/fine_tune_config.json:15-17.

Same constructor, `param_groups` (so LambdaLR / HF schedulers mutate `group["lr"]` as usual), and a
`state_dict()` in torch's layout (`step`, `exp_avg`, `exp_avg_sq` per parameter). The update itself is one
`b2_adamw_grouped` launch per flat buffer instead of ~9 foreach launches per dtype group.
"""
import torch

from . import ops
from ._lib import B2Error
from .flat import flatten_params


class AdamW(torch.optim.Optimizer):
    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=1e-2, amsgrad=False, *,
                 maximize=False, foreach=None, capturable=False, differentiable=False, fused=None,
                 state_dtype=None):
        if isinstance(lr, torch.Tensor):
            raise B2Error("b200_train.AdamW: lr must be a Python float (it is a per-launch scalar)")
        if not 0.0 <= lr:
            raise ValueError(f"Invalid learning rate: {lr}")
        if not 0.0 <= eps:
            raise ValueError(f"Invalid epsilon value: {eps}")
        if not 0.0 <= betas[0] < 1.0:
            raise ValueError(f"Invalid beta parameter at index 0: {betas[0]}")
        if not 0.0 <= betas[1] < 1.0:
            raise ValueError(f"Invalid beta parameter at index 1: {betas[1]}")
        if not 0.0 <= weight_decay:
            raise ValueError(f"Invalid weight_decay value: {weight_decay}")
        if amsgrad or maximize or capturable or differentiable:
            raise B2Error("b200_train.AdamW supports the configuration the reference uses: amsgrad=False, "
                          "maximize=False, capturable=False, differentiable=False")
        # `foreach` / `fused` only choose among torch's own implementations; accepted and ignored.
        defaults = dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay, amsgrad=False, maximize=False,
                        foreach=None, capturable=False, differentiable=False, fused=None,
                        decoupled_weight_decay=True)
        super().__init__(params, defaults)
        if len(self.param_groups) > 8:
            raise B2Error("b200_train.AdamW: at most 8 param_groups")
        self._state_dtype = state_dtype
        self._buffers = None
        self._chunk_group = {}
        self._group_step = []      # one shared CPU step tensor per param_group (torch keeps one per parameter)
        self._fired = set()        # parameters whose gradient was produced by a backward pass since the last step()
        self._skip_tables = {}
        if all(p.is_cuda for g in self.param_groups for p in g["params"]):
            self._setup()           # flatten now, so the very first backward already accumulates into the flat buffer

    # ------------------------------------------------------------------ set-up (first step)
    def _setup(self):
        all_params = [p for g in self.param_groups for p in g["params"]]
        for p in all_params:
            if p.grad is not None and p.grad.is_sparse:
                raise RuntimeError("AdamW does not support sparse gradients")
        self._buffers = flatten_params(all_params)
        gid = {}
        for gi, g in enumerate(self.param_groups):
            for p in g["params"]:
                gid[id(p)] = gi
        self._share_steps()
        for buf in self._buffers:
            sdt = self._state_dtype or buf.dtype      # torch: zeros_like(p) -> state dtype = param dtype
            if buf.exp_avg is None:                   # sharded: state exists for this rank's slices only
                n_state = buf.numel if buf.shards is None else sum(hi - lo for lo, hi, _ in buf.shards)
                buf.exp_avg = torch.zeros(max(n_state, 1), dtype=sdt, device=buf.device)[:n_state]
                buf.exp_avg_sq = torch.zeros(max(n_state, 1), dtype=sdt, device=buf.device)[:n_state]
            table = torch.full((buf.numel // 64,), 255, dtype=torch.uint8)
            for p, o in zip(buf.params, buf.offsets):
                if id(p) in gid:
                    table[o // 64:(o + p.numel() + 63) // 64] = gid[id(p)]
            self._chunk_group[id(buf)] = table.to(buf.device)
            buf.optimizer_attached = True
            for p in buf.params:
                if id(p) in gid and not getattr(p, "_b2_step_hook", False):
                    # the hook is the set's own C-level `add` (no Python frame per parameter on the autograd thread);
                    # the set object therefore has to stay the same one for the optimizer's lifetime (cleared, not replaced)
                    p.register_post_accumulate_grad_hook(self._fired.add)
                    p._b2_step_hook = True
            for p, o in zip(buf.params, buf.offsets):
                if id(p) not in gid:
                    continue
                st = self.state[p]
                if buf.shards is None:
                    # views into the flat state: state_dict() keeps torch's per-parameter layout
                    st["exp_avg"] = buf.exp_avg[o:o + p.numel()].view(p.shape)
                    st["exp_avg_sq"] = buf.exp_avg_sq[o:o + p.numel()].view(p.shape)

    def _table_for(self, buf):
        """Chunk table for this step. torch skips a parameter whose .grad is None (no decay, no state update); here
        gradients are persistent views that are never None, so "received no gradient" is detected instead: if a backward
        pass ran since the last step() and left some parameter untouched, its chunks are marked 255 (= skip). When no
        backward ran at all (gradients written by hand), every parameter is treated as live."""
        base = self._chunk_group[id(buf)]
        if not self._fired:
            return base
        dead = tuple(i for i, p in enumerate(buf.params) if p not in self._fired and p.requires_grad)
        if not dead:
            return base
        key = (id(buf), dead)
        t = self._skip_tables.get(key)
        if t is None:
            host = base.cpu().clone()
            for i in dead:
                o, n = buf.offsets[i], buf.params[i].numel()
                host[o // 64:(o + n + 63) // 64] = 255
            t = host.to(buf.device)
            self._skip_tables[key] = t
        return t

    def _share_steps(self):
        """All parameters of a group advance together, so they share one `step` tensor (saved per parameter by
        state_dict(), as torch does)."""
        self._group_step = []
        for g in self.param_groups:
            cur = 0.0
            for p in g["params"]:
                st = self.state.get(p)
                if st and "step" in st:
                    cur = max(cur, float(st["step"]))
            t = torch.tensor(cur, dtype=torch.float32)
            self._group_step.append(t)
            for p in g["params"]:
                if p.requires_grad:
                    self.state[p]["step"] = t

    def flat_buffers(self):
        if self._buffers is None:
            self._setup()
        return self._buffers

    def add_param_group(self, param_group):
        """Allowed until the flat buffers exist (the constructor goes through here); afterwards the layout of parameters,
        gradients and state is fixed, and a group added later would silently never be updated."""
        if getattr(self, "_buffers", None) is not None:
            raise B2Error("b200_train.AdamW: add_param_group() after the parameters were flattened is not supported; "
                          "pass every param_group to the constructor")
        return super().add_param_group(param_group)

    # ------------------------------------------------------------------ torch.optim.Optimizer surface
    def zero_grad(self, set_to_none=True):
        """ref :273 calls zero_grad(set_to_none=True): the gradients are dropped, as in torch. The next backward pass
        hands each freshly computed gradient to the flat buffer's hook, which copies it into place (flat.FlatBuffer._land)
        — the flat gradient buffer itself is neither cleared nor re-allocated. set_to_none=False zero-fills it instead and
        keeps the `.grad` views."""
        if self._buffers is None:
            return super().zero_grad(set_to_none=set_to_none)
        for buf in self._buffers:
            if set_to_none:
                buf.zero_grad()
            else:
                buf.grad.zero_()
                buf.pending_coef = None
                buf.install_grads()

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        first = self._buffers is None
        if first:
            self._setup()
        for buf in self._buffers:
            buf.install_grads()          # adopts grads autograd allocated before the buffers existed
        hyper = []
        for g, t in zip(self.param_groups, self._group_step):
            t += 1
            b1, b2 = g["betas"]
            hyper.append(dict(lr=g["lr"], beta1=b1, beta2=b2, eps=g["eps"], weight_decay=g["weight_decay"],
                              step=int(t.item())))
        for buf in self._buffers:
            table = self._table_for(buf)
            if buf.shards is None:
                ops.adamw_grouped_(buf.data, buf.grad, buf.exp_avg, buf.exp_avg_sq, table, hyper,
                                   gscale=buf.pending_coef)
            else:
                # this rank's slices only (the ones the reduce-scatter completed), front of the buffer first; after each
                # one the wrapper pushes that range of the updated parameters to every peer (on its side stream)
                for k, (lo, hi, so) in enumerate(buf.shards):
                    if hi > lo:
                        ops.adamw_grouped_(buf.data[lo:hi], buf.grad[lo:hi], buf.exp_avg[so:so + hi - lo],
                                           buf.exp_avg_sq[so:so + hi - lo], table[lo // 64:hi // 64], hyper,
                                           gscale=buf.pending_coef)
                    if buf.after_shard_update is not None:
                        buf.after_shard_update(k)
            buf.pending_coef = None
        self._fired.clear()
        return loss

    def _gathered_state(self, buf):
        """Full-length exp_avg / exp_avg_sq of a sharded buffer (checkpoint path only; one NCCL all-reduce each: every
        element is owned by exactly one rank, the others contribute zeros)."""
        import torch.distributed as dist
        full = []
        for shard in (buf.exp_avg, buf.exp_avg_sq):
            out = torch.zeros(buf.numel, dtype=shard.dtype, device=shard.device)
            for lo, hi, so in buf.shards:
                out[lo:hi] = shard[so:so + hi - lo]
            dist.all_reduce(out, group=buf.comm.group)
            full.append(out)
        return full

    def state_dict(self):
        """torch's layout (`step`, `exp_avg`, `exp_avg_sq` per parameter, ref: pytorch_llm_ray.py:302) even when the
        state is sharded over ranks: the shards are gathered for the occasion. Collective: call it on every rank."""
        sharded = [b for b in (self._buffers or []) if b.shards is not None]
        for buf in sharded:
            m, v = self._gathered_state(buf)
            for p, o in zip(buf.params, buf.offsets):
                if p in self.state:
                    self.state[p]["exp_avg"] = m[o:o + p.numel()].view(p.shape)
                    self.state[p]["exp_avg_sq"] = v[o:o + p.numel()].view(p.shape)
        sd = super().state_dict()
        # Detach the per-parameter dicts from self.state, and give every parameter its OWN step tensor: internally one
        # tensor per group is shared, but torch's optimizers increment each entry they are handed (foreach), so a
        # checkpoint carrying a shared tensor would advance several steps at once after being loaded there.
        sd["state"] = {k: dict(v) for k, v in sd["state"].items()}
        for v in sd["state"].values():
            if isinstance(v.get("step"), torch.Tensor):
                v["step"] = v["step"].clone()
        for buf in sharded:
            for p in buf.params:
                if p in self.state:
                    self.state[p].pop("exp_avg", None)
                    self.state[p].pop("exp_avg_sq", None)
        return sd

    def load_state_dict(self, state_dict):
        super().load_state_dict(state_dict)
        if self._buffers is None:
            return
        # re-home the loaded tensors into the flat state so the kernels see them
        for buf in self._buffers:
            segs = buf.shards if buf.shards is not None else [(0, buf.numel, 0)]
            for p, o in zip(buf.params, buf.offsets):
                st = self.state.get(p)
                if not st or "exp_avg" not in st:
                    continue
                for key, flat in (("exp_avg", buf.exp_avg), ("exp_avg_sq", buf.exp_avg_sq)):
                    for lo, hi, so in segs:
                        a, b = max(o, lo), min(o + p.numel(), hi)      # part of this parameter that lives in our state
                        if a < b:
                            flat[so + a - lo:so + b - lo].copy_(st[key].reshape(-1)[a - o:b - o])
                    if buf.shards is None:
                        st[key] = flat[o:o + p.numel()].view(p.shape)
                    else:
                        st.pop(key)
        self._share_steps()
