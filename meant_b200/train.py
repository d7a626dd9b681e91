"""Minimal data-parallel training step for the MEANT models (the reference's loop is
src/in_loop_train.py:196-330: autocast forward, CrossEntropyLoss on the sigmoid outputs,
backward, clip_grad_norm_(1.0), AdamW step).  bf16 autocast replaces fp16 + GradScaler."""
import torch
import torch.distributed as dist
import torch.nn.functional as F

from .ddp import BucketedGradAllReduce, broadcast_parameters
from .optim import FusedAdamW, prob_cross_entropy


class TrainStep:
    """graph=True (CUDA): after `graph_warmup` eager steps the whole step -- zero_grad, autocast forward, loss,
    backward, the bucketed NCCL gradient all-reduces (captured on NCCL's stream, still overlapping backward),
    clip, AdamW -- is captured once into a CUDA graph and replayed, so the ~2000 launches of a step cost no
    host time; inputs are copied into static buffers, dropout masks advance through
    a device-resident seed, and the bias-gradient kernels replace autograd's reduction.  All of that state (weight
    shadows, the zero arena, the seed, the Linear routing) lives in this instance's own ops.OpsContext (`self.ctx`),
    active only around its forward + backward: two TrainSteps / models never see each other's.  Input shapes and non-tensor arguments must not change between calls."""

    def __init__(self, model, lr=5e-5, betas=(0.9, 0.999), weight_decay=0.0, max_grad_norm=1.0,
                 bucket_mb=25.0, bf16=True, group=None, graph=False, graph_warmup=3, own_optimizer=True, grad_dtype=None):
        self.model = model
        self.bf16 = bf16
        self.max_grad_norm = max_grad_norm
        broadcast_parameters(model, 0, group)
        self.params = [p for p in model.parameters() if p.requires_grad]
        fused = self._fused = self.params[0].is_cuda
        own = bool(own_optimizer) and fused and all(p.dtype == torch.float32 and p.is_contiguous() for p in self.params)
        if grad_dtype is not None and not own:
            raise ValueError("grad_dtype (compressed gradient buckets) needs the library's optimizer (CUDA, fp32 parameters)")
        self.reducer = BucketedGradAllReduce(self.params, bucket_mb=bucket_mb, group=group, grad_dtype=grad_dtype)
        self.graph = bool(graph) and fused
        self._graph_warmup, self._calls = int(graph_warmup), 0
        self._g = self._static = self._loss = self._seed_t = None
        self.launches_per_replay = 0            # native-library launches captured in one step
        self._shadow_src = self._shadow_dst = self._shadows = self._arena = None
        from . import ops
        if self.graph:
            if bf16:
                self._build_shadows()
            self._arena = torch.zeros(1 << 20, dtype=torch.float32, device=self.params[0].device)
        self.ctx = ops.OpsContext(shadow=self._shadows, arena=self._arena, fused_linear=self.graph)
        # The step after backward on the library's multi-tensor kernels (optim.py: gradient norm, clip + AdamW, the bf16
        # weight shadows written by the update itself) -- or torch's optimizer (CPU tensors: the gloo tests)
        self._own = own
        if self._own:
            self.opt = FusedAdamW(self.params, lr=lr, betas=betas, weight_decay=weight_decay, max_norm=max_grad_norm or 0.0,
                                  shadows=self._shadows)
            self.refresh_shadows()
        else:
            self.opt = torch.optim.AdamW(self.params, lr=lr, betas=betas, weight_decay=weight_decay, fused=fused,
                                         capturable=self.graph)

    def refresh_shadows(self):
        """Re-derive the bf16 weight shadows from the fp32 parameters.  With the library's optimizer the update keeps
        them current, so this runs once here -- call it again after changing parameters from outside (load_state_dict)."""
        if self._shadow_src:
            torch._foreach_copy_(self._shadow_dst, self._shadow_src)

    def _build_shadows(self):
        """bf16 copies of every nn.Linear weight / bias, refreshed by ONE multi-tensor cast per step; the q, v, k
        shadows of an attention module are row blocks of one packed buffer (ops.linear3 uses it as is)."""
        import torch.nn as nn
        shadows, src, dst, seen = {}, [], [], set()

        def add(p, view):
            shadows[id(p)] = (p, view)
            src.append(p.detach())
            dst.append(view)
            seen.add(id(p))

        for m in self.model.modules():
            trio = [getattr(m, n, None) for n in ("q", "v", "k")]
            if all(isinstance(t, nn.Linear) and t.bias is not None and t.weight.is_cuda for t in trio) and \
                    len({t.weight.shape[1] for t in trio}) == 1 and not any(id(t.weight) in seen for t in trio):
                rows = [t.weight.shape[0] for t in trio]
                pw = torch.empty(sum(rows), trio[0].weight.shape[1], dtype=torch.bfloat16, device=trio[0].weight.device)
                pb = torch.empty(sum(rows), dtype=torch.bfloat16, device=pw.device)
                off = 0
                for t, r in zip(trio, rows):
                    add(t.weight, pw[off:off + r])
                    add(t.bias, pb[off:off + r])
                    off += r
                shadows[("packed",) + tuple(id(t.weight) for t in trio)] = (pw, pb, [t.weight for t in trio])
        for m in self.model.modules():
            if isinstance(m, nn.Linear) and m.weight.is_cuda and m.weight.dtype == torch.float32:
                for p in (m.weight, m.bias):
                    if p is not None and id(p) not in seen:
                        add(p, torch.empty_like(p, dtype=torch.bfloat16))
        self._shadows, self._shadow_src, self._shadow_dst = shadows, src, dst

    def _step(self, labels, args, kwargs):
        from . import ops
        if self._shadow_src and not self._own:
            torch._foreach_copy_(self._shadow_dst, self._shadow_src)
        if self._arena is not None:
            self._arena.zero_()
        self.ctx.begin_step()
        with ops.use(self.ctx):
            return self._step_body(labels, args, kwargs)

    def _step_body(self, labels, args, kwargs):
        self.reducer.zero_grad()
        device_type = "cuda" if self.params[0].is_cuda else "cpu"
        with torch.autocast(device_type, dtype=torch.bfloat16, enabled=self.bf16 and device_type == "cuda"):
            out = self.model(*args, **kwargs)
        loss = prob_cross_entropy(out, labels) if self._own else F.cross_entropy(out.float(), labels)
        loss.backward()
        self.reducer.finish()
        if self._own:
            self.opt.step(self.reducer.grads())  # gradient norm + clip + AdamW (+ shadows): optim.py
            return loss.detach()
        if self.max_grad_norm and self._fused:
            # clip_grad_norm_ without its own pass over the gradients: fused AdamW divides every gradient by
            # `grad_scale` (and writes it back) inside its update kernel, so the clip coefficient rides along
            grads = [p.grad for p in self.params if p.grad is not None]
            total = torch.linalg.vector_norm(torch.stack(torch._foreach_norm(grads)))
            self.opt.grad_scale = torch.clamp((total + 1e-6) / self.max_grad_norm, min=1.0).float().reshape(())
        elif self.max_grad_norm:
            torch.nn.utils.clip_grad_norm_(self.params, self.max_grad_norm, foreach=True)
        self.opt.step()
        return loss.detach()

    def _capture(self, labels, args, kwargs):
        from . import ops, _native as nv
        clone = lambda v: v.clone() if torch.is_tensor(v) else v
        self._static = (clone(labels), tuple(clone(a) for a in args), {k: clone(v) for k, v in kwargs.items()})
        self._seed_t = torch.zeros(1, dtype=torch.int64, device=labels.device)
        torch.cuda.synchronize()
        self._g = torch.cuda.CUDAGraph()
        self.ctx.graph_seed = self._seed_t
        n0 = nv.launch_count()
        try:
            with torch.cuda.graph(self._g):
                self._seed_t.add_(0x1E3779B97F4A7C15)     # every replay shifts all dropout seeds
                self._loss = self._step(self._static[0], self._static[1], self._static[2])
        finally:
            self.ctx.graph_seed = None
        self.launches_per_replay = nv.launch_count() - n0

    def _load(self, labels, args, kwargs):
        s_labels, s_args, s_kwargs = self._static
        pairs = [(s_labels, labels)] + list(zip(s_args, args)) + [(s_kwargs[k], kwargs[k]) for k in s_kwargs]
        if len(args) != len(s_args) or set(kwargs) != set(s_kwargs):
            raise ValueError("TrainStep(graph=True): the call signature changed after capture")
        for dst, src in pairs:
            if torch.is_tensor(dst):
                if not torch.is_tensor(src) or src.shape != dst.shape or src.dtype != dst.dtype:
                    raise ValueError("TrainStep(graph=True): input shapes / dtypes are fixed after capture")
                if src.data_ptr() != dst.data_ptr():
                    dst.copy_(src, non_blocking=True)
            elif dst is not src and dst != src:
                raise ValueError("TrainStep(graph=True): non-tensor arguments are fixed after capture")

    def close(self):
        """Release the captured graph (NCCL communicators cannot be destroyed while a graph holds their kernels)."""
        if self._g is not None:
            torch.cuda.synchronize()
            self._g.reset()
            self._g, self._calls = None, 0
            self._static = self._loss = None
            self.reducer._handles = []          # work handles created under capture never complete outside the graph

    def __call__(self, labels, *args, **kwargs):
        """One optimisation step; returns the (device) loss tensor (graph mode: a static tensor the next
        call overwrites)."""
        self._calls += 1
        if not self.graph or self._calls <= self._graph_warmup:
            return self._step(labels, args, kwargs)
        if self._g is None:
            self._capture(labels, args, kwargs)
        else:
            self._load(labels, args, kwargs)
        self._g.replay()
        return self._loss
