"""meant_vqa (no lag axis) and the MLM pre-trainer stack.
Reference: meant/meant_vqa.py:143-233, pretrain_mlm.py:74-88."""
import torch
from torch import nn

from . import ops
from .meant import languageEncoder, visionEncoder, _PatchRearrange, _run_stack
from .rms_norm import RMSNorm


class meant_vqa(nn.Module):
    def __init__(self, text_dim, image_dim, price_dim, height, width, patch_res, lag, num_classes, embedding,
                 flash=False, num_heads=8, num_encoders=1, channels=4):
        super().__init__()
        self.lag, self.text_dim, self.image_dim = lag, text_dim, image_dim
        self.dim = text_dim + image_dim
        self.num_heads, self.channels = num_heads, channels
        self.patch_dim = channels * patch_res * patch_res
        self.n = int((height * width) / (patch_res ** 2))
        self.embedding = nn.ModuleList([embedding])
        self.patchEmbed = nn.Sequential(_PatchRearrange(patch_res), nn.Linear(self.patch_dim, image_dim))
        self.visionEncoders = nn.ModuleList([visionEncoder(image_dim, num_heads, flash=flash)
                                             for _ in range(num_encoders)])
        self.languageEncoders = nn.ModuleList([languageEncoder(text_dim, num_heads, flash=flash)
                                               for _ in range(num_encoders)])
        # constructed but unused by the reference forward (meant/meant_vqa.py:199-200,225-229);
        # kept because they are part of the state_dict
        self.multimodal_embedding = nn.Sequential(nn.Linear(1, self.dim), nn.GELU(), RMSNorm(self.dim),
                                                  nn.Linear(self.dim, self.dim))
        self.multimodal_encoding = nn.ModuleList([visionEncoder(self.dim, num_heads, flash=flash)])
        self.mlpHead = nn.ModuleList([RMSNorm(self.dim), nn.Linear(self.dim, num_classes), nn.Sigmoid()])

    def forward(self, tweets=None, images=None, attention_mask=None, **kw):
        words = kw.get("input_ids", tweets)
        images = kw.get("pixels", images)
        for mod in self.embedding:
            words = ops.embed(mod, words)
        words = _run_stack(self.languageEncoders, words, attention_mask)
        img = _run_stack(self.visionEncoders, self.patchEmbed(images))
        t = torch.cat((ops.mean_pool(words), ops.mean_pool(img)), dim=1)
        for mod in self.mlpHead:
            t = mod(t)
        return t.float()


class meant_language_pretrainer(nn.Module):
    def __init__(self, num_encoders, mlm_input_dim, embedding, lm_head, flash=False, lag=5, text_dim=768,
                 num_heads=8):
        super().__init__()
        self.embedding = nn.ModuleList([embedding])
        self.languageEncoders = nn.ModuleList([languageEncoder(text_dim, num_heads, flash=flash)
                                               for _ in range(num_encoders)])
        self.mlm_head = lm_head
        self.lag = lag

    def forward(self, words, attention_mask=None):
        for mod in self.embedding:
            words = ops.embed(mod, words)
        return self.mlm_head(_run_stack(self.languageEncoders, words, attention_mask))
