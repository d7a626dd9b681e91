"""meant_vision: image-only MEANT over the lag window.  Reference: meant/meant_vision.py:107-165."""
from torch import nn

from . import ops
from .meant import visionEncoder, _PatchRearrange, _run_stack
from .meant_tweet import temporalEncoder, _LayerNorm


class meant_vision(nn.Module):
    def __init__(self, image_dim, price_dim, height, width, patch_res, lag, num_classes, flash=False, num_heads=8,
                 num_encoders=1, channels=4):
        super().__init__()
        self.dim, self.num_heads, self.channels = image_dim, num_heads, channels
        self.patch_dim = channels * patch_res * patch_res
        self.n = int((height * width) / (patch_res ** 2))
        self.patchEmbed = nn.Sequential(_PatchRearrange(patch_res), nn.Linear(self.patch_dim, image_dim))
        self.visionEncoders = nn.ModuleList([visionEncoder(image_dim, num_heads, flash=flash)
                                             for _ in range(num_encoders)])
        self.temporal_encoding = nn.ModuleList([temporalEncoder(self.dim, num_heads, lag)])
        self.mlpHead = nn.ModuleList([_LayerNorm(self.dim), nn.Linear(self.dim, num_classes), nn.Sigmoid()])

    def forward(self, images=None, prices=None, **kw):
        images = kw.get("pixels", images)
        b, lag = images.shape[:2]
        img = self.patchEmbed(images.reshape(b * lag, *images.shape[2:]))
        img = _run_stack(self.visionEncoders, img)
        t = ops.mean_pool(img).view(b, lag, -1)
        for enc in self.temporal_encoding:
            t = enc(t)
        for mod in self.mlpHead:
            t = mod(t)
        return t.squeeze(dim=1).float()
