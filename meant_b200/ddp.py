"""Data-parallel gradient exchange for the MEANT training step: one process per GPU, flat
gradient buckets all-reduced (NCCL over NVLink 5 / NVSwitch) as soon as backward has produced
them, overlapping the rest of backward.  The reference has no distributed code at all
(only nn.DataParallel in pretrain_mlm.py:330); this is the B200-native replacement.

Design
  * parameters are walked in REVERSE registration order (head -> temporal -> encoders ->
    embedding: the order backward produces gradients) and packed into ~bucket_mb buckets;
  * gradients are produced by autograd as usual (.grad starts the step as None, so the first
    arrival is a move, not a zero-fill plus an add per parameter); a post-accumulate-grad hook
    counts arrivals, and the last arrival of a bucket packs the bucket's gradients into its flat
    buffer with ONE multi-tensor copy, re-points every .grad at its slice of the buffer and
    launches all_reduce(bucket, AVG, async_op=True) -- NCCL runs it on its own stream;
  * with one rank nothing is packed at all: the optimizer reads the autograd tensors;
  * finish() waits for the outstanding handles before the optimizer reads the gradients;
  * grad_dtype=torch.bfloat16 exchanges bf16 buckets (half the bytes on the wire): gradients are cast while they are
    packed, all-reduced in bf16, and handed to the optimizer as bf16 slices through grads() -- .grad is dropped then,
    because torch does not allow a bf16 .grad on an fp32 parameter (meant_b200.optim.FusedAdamW reads either type).
The same code runs on gloo (CPU tensors) for the world_size-2 tests.
"""
import torch
import torch.distributed as dist


class BucketedGradAllReduce:
    def __init__(self, params, bucket_mb: float = 25.0, group=None, grad_dtype=None):
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_available() and dist.is_initialized() else 1
        self.params = [p for p in params if p.requires_grad]
        self.grad_dtype = grad_dtype
        self._reduced = {}         # compressed mode: param -> reduced bucket slice of this step
        self.buckets = []          # [(flat tensor, [params])]
        self._bucket_of = {}
        self._pending = []
        self._handles = []
        self._views, self._numel = [], []
        cap = int(bucket_mb * 1024 * 1024)
        cur, cur_bytes = [], 0
        for p in reversed(self.params):
            nbytes = p.numel() * (torch.empty((), dtype=grad_dtype or p.dtype).element_size())
            if cur and cur_bytes + nbytes > cap:
                self._seal(cur, grad_dtype)
                cur, cur_bytes = [], 0
            cur.append(p)
            cur_bytes += nbytes
        if cur:
            self._seal(cur, grad_dtype)
        self._pending = [len(ps) for _, ps in self.buckets]
        self._next = 0             # first bucket not yet handed to the communicator this step
        self._hooks = [p.register_post_accumulate_grad_hook(self._on_grad) for p in self.params]

    def _seal(self, ps, grad_dtype):
        dtype = grad_dtype or ps[0].dtype
        flat = torch.empty(sum(p.numel() for p in ps), dtype=dtype, device=ps[0].device) if self.world > 1 else None
        views, off = [], 0
        for p in ps:
            if flat is not None:
                views.append(flat[off:off + p.numel()].view_as(p))
            off += p.numel()
            self._bucket_of[p] = len(self.buckets)
        self._views.append(views)
        self._numel.append(off)
        self.buckets.append((flat, ps))

    def zero_grad(self):
        """Drop the gradients (autograd then moves instead of accumulating) and re-arm the counters."""
        for p in self.params:
            p.grad = None
        self._pending = [len(ps) for _, ps in self.buckets]
        self._handles = []
        self._next = 0
        self._reduced = {}

    def grads(self):
        """{param: gradient} for the optimizer after finish(): the reduced bucket slices (bf16 in compressed mode), or
        with one rank the autograd tensors."""
        if self._reduced:
            return dict(self._reduced)
        return {p: p.grad for p in self.params if p.grad is not None}

    def _reduce(self, b):
        """Pack bucket b and all-reduce it.  A parameter without a local gradient this step contributes zeros
        -- the correct term of the average when another rank did produce one -- and EVERY parameter of the
        bucket ends up with .grad = its reduced slice, so all ranks hand the optimizer the same gradients and the
        replicas cannot drift apart (a parameter no rank used gets an all-zero gradient, as under
        torch's DistributedDataParallel)."""
        flat, ps = self.buckets[b]
        views = self._views[b]
        have = [(v, p) for v, p in zip(views, ps) if p.grad is not None]
        if len(have) != len(ps):
            flat.zero_()
        if have:
            torch._foreach_copy_([v for v, _ in have], [p.grad for _, p in have])
        compressed = self.grad_dtype is not None and self.grad_dtype != ps[0].dtype
        for v, p in zip(views, ps):
            if compressed:
                self._reduced[p] = v
                p.grad = None
            else:
                p.grad = v
        if flat.is_cuda:
            self._handles.append(dist.all_reduce(flat, op=dist.ReduceOp.AVG, group=self.group, async_op=True))
        else:  # gloo has no AVG
            flat.div_(self.world)
            self._handles.append(dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=self.group, async_op=True))

    def _on_grad(self, p):
        b = self._bucket_of[p]
        self._pending[b] -= 1
        if self._pending[b] < 0:
            # a second backward() before zero_grad(): the bucket was already reduced, the new micro-batch would
            # accumulate locally and never be exchanged
            raise RuntimeError("BucketedGradAllReduce: a parameter received a second gradient before zero_grad(); "
                               "gradient accumulation over several backward passes is not supported -- call "
                               "zero_grad() between them")
        if self._pending[b] == 0 and self.world > 1:
            # collectives are issued in bucket order on every rank (a bucket that completes early waits for its
            # predecessors), so a rank on which some parameter got no gradient cannot reorder them
            while self._next < len(self.buckets) and self._pending[self._next] == 0:
                self._reduce(self._next)
                self._next += 1

    def finish(self):
        """Wait for every bucket; buckets some of whose parameters got no gradient this step are reduced
        here so that all ranks issue the same collectives."""
        if self.world > 1:
            while self._next < len(self.buckets):
                self._reduce(self._next)
                self._pending[self._next] = 0
                self._next += 1
        for h in self._handles:
            h.wait()
        self._handles = []

    def bucket_bytes(self):
        return [n * ps[0].element_size() for n, (_, ps) in zip(self._numel, self.buckets)]


def broadcast_parameters(module, src=0, group=None):
    """Make every rank start from rank `src`'s weights (what DDP's constructor does)."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return
    for t in list(module.parameters()) + list(module.buffers()):
        dist.broadcast(t.data, src=src, group=group)
