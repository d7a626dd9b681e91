"""Minimal data-parallel training step for the MEANT models (the reference's loop is
src/in_loop_train.py:196-330: autocast forward, CrossEntropyLoss on the sigmoid outputs,
backward, clip_grad_norm_(1.0), AdamW step).  bf16 autocast replaces fp16 + GradScaler."""
import torch
import torch.distributed as dist
import torch.nn.functional as F

from .ddp import BucketedGradAllReduce, broadcast_parameters


class TrainStep:
    def __init__(self, model, lr=5e-5, betas=(0.9, 0.999), weight_decay=0.0, max_grad_norm=1.0,
                 bucket_mb=25.0, bf16=True, group=None):
        self.model = model
        self.bf16 = bf16
        self.max_grad_norm = max_grad_norm
        broadcast_parameters(model, 0, group)
        self.params = [p for p in model.parameters() if p.requires_grad]
        self.reducer = BucketedGradAllReduce(self.params, bucket_mb=bucket_mb, group=group)
        fused = self.params[0].is_cuda
        self.opt = torch.optim.AdamW(self.params, lr=lr, betas=betas, weight_decay=weight_decay, fused=fused)

    def __call__(self, labels, *args, **kwargs):
        """One optimisation step; returns the (device) loss tensor."""
        self.reducer.zero_grad()
        device_type = "cuda" if self.params[0].is_cuda else "cpu"
        with torch.autocast(device_type, dtype=torch.bfloat16, enabled=self.bf16 and device_type == "cuda"):
            out = self.model(*args, **kwargs)
        loss = F.cross_entropy(out.float(), labels)
        loss.backward()
        self.reducer.finish()
        if self.max_grad_norm:
            torch.nn.utils.clip_grad_norm_(self.params, self.max_grad_norm, foreach=True)
        self.opt.step()
        return loss.detach()
