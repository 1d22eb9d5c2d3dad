"""ViT-B/16 callers of the hot path (BASELINE.json configs[3]).

Mirrors models/vit_quantized_approx.py:56-394 -- the same wrapper classes, the same places where
activations are re-quantized -- over an FP32 graph with the layout of the HuggingFace ViT the reference
wraps (`ViTForImageClassification`, config defaults: 12 layers, hidden 768, MLP 3072, 12 heads, 16x16
patches, 224x224 images -> 197 tokens).  The FP32 graph is defined here instead of imported: the
reference names `ViTSdpaSelfAttention` / `ViTSdpaAttention`, which transformers 5.5 no longer has
(SURVEY.md 8(c)), and there is no network for the pretrained checkpoint, so weights are seeded random.

Hot-path content per image: 12 x (4 x 197x768x768 + 197x768x3072 + 197x3072x768) approx MACs through
QCustomLinearTorch (3-D activations are flattened to [B*197, K]: the reference asserts on them,
SURVEY.md 0.9); the patch-embedding conv has no BatchNorm -> QCustomConv2dTorch -> exact matmul
(tensor cores); attention scores/values stay inside scaled_dot_product_attention, not quantized.
"""
import torch
from torch import nn
import torch.nn.functional as F

from .model_wrap import QuantLayerNorm, QuantizedModel, quantize_model   # noqa: F401  (QuantLayerNorm re-exported)
from .quantization import QuantizedActivation


# ---- FP32 graph (HuggingFace ViT layout) -----------------------------------------------------------
class ViTConfig:
    def __init__(self, hidden_size=768, num_hidden_layers=12, num_attention_heads=12, intermediate_size=3072,
                 image_size=224, patch_size=16, num_channels=3, layer_norm_eps=1e-12, num_labels=1000):
        self.__dict__.update(locals())
        del self.__dict__["self"]


class ViTPatchEmbeddings(nn.Module):
    def __init__(self, c):
        super().__init__()
        self.num_channels, self.image_size, self.patch_size = c.num_channels, (c.image_size,) * 2, (c.patch_size,) * 2
        self.num_patches = (c.image_size // c.patch_size) ** 2
        self.projection = nn.Conv2d(c.num_channels, c.hidden_size, c.patch_size, c.patch_size)

    def forward(self, x):
        return self.projection(x).flatten(2).transpose(1, 2)


class ViTEmbeddings(nn.Module):
    def __init__(self, c):
        super().__init__()
        self.patch_embeddings = ViTPatchEmbeddings(c)
        self.cls_token = nn.Parameter(torch.randn(1, 1, c.hidden_size) * 0.02)
        self.position_embeddings = nn.Parameter(torch.randn(1, self.patch_embeddings.num_patches + 1, c.hidden_size) * 0.02)
        self.dropout = nn.Dropout(0.0)

    def forward(self, x):
        e = self.patch_embeddings(x)
        e = torch.cat((self.cls_token.expand(x.shape[0], -1, -1), e), dim=1)
        return self.dropout(e + self.position_embeddings)


class ViTSelfAttention(nn.Module):
    def __init__(self, c):
        super().__init__()
        self.num_attention_heads = c.num_attention_heads
        self.attention_head_size = c.hidden_size // c.num_attention_heads
        self.all_head_size = c.hidden_size
        self.query, self.key, self.value = (nn.Linear(c.hidden_size, c.hidden_size) for _ in range(3))
        self.dropout = nn.Dropout(0.0)
        self.attention_probs_dropout_prob = 0.0

    def transpose_for_scores(self, x):
        return x.view(x.size()[:-1] + (self.num_attention_heads, self.attention_head_size)).permute(0, 2, 1, 3)

    def forward(self, x):
        q, k, v = (self.transpose_for_scores(f(x)) for f in (self.query, self.key, self.value))
        ctx = F.scaled_dot_product_attention(q, k, v).permute(0, 2, 1, 3).contiguous()
        return ctx.view(ctx.size()[:-2] + (self.all_head_size,))


class ViTSelfOutput(nn.Module):
    def __init__(self, c):
        super().__init__()
        self.dense = nn.Linear(c.hidden_size, c.hidden_size)
        self.dropout = nn.Dropout(0.0)

    def forward(self, x):
        return self.dropout(self.dense(x))


class ViTAttention(nn.Module):
    def __init__(self, c):
        super().__init__()
        self.attention, self.output = ViTSelfAttention(c), ViTSelfOutput(c)

    def forward(self, x):
        return self.output(self.attention(x))


class ViTIntermediate(nn.Module):
    def __init__(self, c):
        super().__init__()
        self.dense = nn.Linear(c.hidden_size, c.intermediate_size)
        self.intermediate_act_fn = nn.GELU()

    def forward(self, x):
        return self.intermediate_act_fn(self.dense(x))


class ViTOutput(nn.Module):
    def __init__(self, c):
        super().__init__()
        self.dense = nn.Linear(c.intermediate_size, c.hidden_size)
        self.dropout = nn.Dropout(0.0)

    def forward(self, x, inp):
        return self.dropout(self.dense(x)) + inp


class ViTLayer(nn.Module):
    def __init__(self, c):
        super().__init__()
        self.attention, self.intermediate, self.output = ViTAttention(c), ViTIntermediate(c), ViTOutput(c)
        self.layernorm_before = nn.LayerNorm(c.hidden_size, eps=c.layer_norm_eps)
        self.layernorm_after = nn.LayerNorm(c.hidden_size, eps=c.layer_norm_eps)

    def forward(self, h):
        h = self.attention(self.layernorm_before(h)) + h
        return self.output(self.intermediate(self.layernorm_after(h)), h)


class ViTEncoder(nn.Module):
    def __init__(self, c):
        super().__init__()
        self.layer = nn.ModuleList(ViTLayer(c) for _ in range(c.num_hidden_layers))
        self.gradient_checkpointing = False

    def forward(self, h):
        for layer in self.layer:
            h = layer(h)
        return h


class ViTModel(nn.Module):
    def __init__(self, c):
        super().__init__()
        self.embeddings, self.encoder = ViTEmbeddings(c), ViTEncoder(c)
        self.layernorm = nn.LayerNorm(c.hidden_size, eps=c.layer_norm_eps)

    def forward(self, x):
        return self.layernorm(self.encoder(self.embeddings(x)))


class VisionTransformerForImageClassification(nn.Module):
    def __init__(self, config=None):
        super().__init__()
        self.config = config or ViTConfig()
        self.vit = ViTModel(self.config)
        self.classifier = nn.Linear(self.config.hidden_size, self.config.num_labels)
        for m in self.modules():                     # HF _init_weights: N(0, 0.02), zero bias
            if isinstance(m, (nn.Linear, nn.Conv2d)):
                m.weight.data.normal_(0.0, 0.02)
                m.bias.data.zero_()

    def forward(self, x):
        return self.classifier(self.vit(x)[:, 0, :])


# ---- quantized wrappers (models/vit_quantized_approx.py:56-394) --------------------------------------
def _q(module, specials=None, **quant_params):
    """quantize_model over the reference's non-BN map (Conv2d, Linear, LayerNorm -> QuantLayerNorm,
    approx/replace_operations_with_approx_ops.py:180-187)."""
    return quantize_model(module, specials=specials or {}, **quant_params)


class QuantizedVitPatchEmbeddings(QuantizedActivation):
    def __init__(self, orig, **quant_params):
        super().__init__(**quant_params)
        self.projection = _q(orig.projection, **quant_params)       # conv without BN: exact matmul (tensor cores)
        self.num_channels, self.image_size = orig.num_channels, orig.image_size
        self.patch_size, self.num_patches = orig.patch_size, orig.num_patches

    def forward(self, pixel_values):
        _, c, h, w = pixel_values.shape
        if c != self.num_channels:
            raise ValueError("Make sure that the channel dimension of the pixel values match with the one set in "
                             f"the configuration. Expected {self.num_channels} but got {c}.")
        if h != self.image_size[0] or w != self.image_size[1]:
            raise ValueError(f"Input image size ({h}*{w}) doesn't match model ({self.image_size[0]}*{self.image_size[1]}).")
        return self.quantize_activations(self.projection(pixel_values).flatten(2).transpose(1, 2))


class QuantizedVitEmbeddings(QuantizedActivation):
    def __init__(self, orig, **quant_params):
        super().__init__(**quant_params)
        self.patch_embeddings = QuantizedVitPatchEmbeddings(orig.patch_embeddings, **quant_params)
        self.cls_token, self.position_embeddings, self.dropout = orig.cls_token, orig.position_embeddings, orig.dropout

    def forward(self, x):
        e = self.patch_embeddings(x)
        e = torch.cat((self.cls_token.expand(x.shape[0], -1, -1), e), dim=1) + self.position_embeddings
        return self.quantize_activations(self.dropout(e))


class QuantizedViTImmediate(QuantizedActivation):
    def __init__(self, orig, **quant_params):
        super().__init__(**quant_params)
        self.dense = _q(orig.dense, **quant_params)
        self.intermediate_act_fn = orig.intermediate_act_fn

    def forward(self, h):
        return self.quantize_activations(self.intermediate_act_fn(self.dense(h)))


class QuantizedViTOutput(QuantizedActivation):
    def __init__(self, orig, **quant_params):
        super().__init__(**quant_params)
        self.dense, self.dropout = _q(orig.dense, **quant_params), orig.dropout

    def forward(self, h, input_tensor):
        return self.quantize_activations(self.dropout(self.dense(h)) + input_tensor)


class QuantizedViTSelfAttention(QuantizedActivation):
    def __init__(self, orig, **quant_params):
        super().__init__(**quant_params)
        self.num_attention_heads, self.attention_head_size = orig.num_attention_heads, orig.attention_head_size
        self.all_head_size = orig.all_head_size
        self.query, self.key, self.value = (_q(m, **quant_params) for m in (orig.query, orig.key, orig.value))
        self.dropout = orig.dropout
        self.attention_probs_dropout_prob = orig.attention_probs_dropout_prob

    def transpose_for_scores(self, x):
        return x.view(x.size()[:-1] + (self.num_attention_heads, self.attention_head_size)).permute(0, 2, 1, 3)

    def forward(self, x):
        q = self.query(x)
        k, v = self.transpose_for_scores(self.key(x)), self.transpose_for_scores(self.value(x))
        ctx = F.scaled_dot_product_attention(self.transpose_for_scores(q), k, v, attn_mask=None,
                                             dropout_p=self.attention_probs_dropout_prob if self.training else 0.0,
                                             is_causal=False, scale=None)
        ctx = ctx.permute(0, 2, 1, 3).contiguous()
        return self.quantize_activations(ctx.view(ctx.size()[:-2] + (self.all_head_size,)))


class QuantizedViTSelfOutput(QuantizedActivation):
    def __init__(self, orig, **quant_params):
        super().__init__(**quant_params)
        self.dense, self.dropout = _q(orig.dense, **quant_params), orig.dropout

    def forward(self, x):
        return self.dropout(self.dense(x))


class QuantizedViTSdpaAttention(QuantizedActivation):
    def __init__(self, orig, **quant_params):
        super().__init__(**quant_params)
        self.attention = QuantizedViTSelfAttention(orig.attention, **quant_params)
        self.output = QuantizedViTSelfOutput(orig.output, **quant_params)

    def forward(self, x):
        return self.output(self.attention(x))


class QuantizedViTLayer(QuantizedActivation):
    def __init__(self, orig, **quant_params):
        super().__init__(**quant_params)
        self.intermediate = QuantizedViTImmediate(orig.intermediate, **quant_params)
        self.attention = QuantizedViTSdpaAttention(orig.attention, **quant_params)
        self.output = QuantizedViTOutput(orig.output, **quant_params)
        self.layernorm_before = _q(orig.layernorm_before, **quant_params)
        self.layernorm_after = _q(orig.layernorm_after, **quant_params)

    def forward(self, hidden_states):
        attention_output = self.attention(self.layernorm_before(hidden_states))
        hidden_states = self.quantize_activations(attention_output + hidden_states)      # first residual
        layer_output = self.intermediate(self.layernorm_after(hidden_states))
        return self.output(layer_output, hidden_states)                                 # second residual, quantized there


class QuantizedViTEncoder(QuantizedActivation):
    def __init__(self, orig, **quant_params):
        super().__init__(**quant_params)
        self.layer = nn.ModuleList(QuantizedViTLayer(m, **quant_params) for m in orig.layer)
        self.gradient_checkpointing = orig.gradient_checkpointing

    def forward(self, hidden_states):
        for layer_module in self.layer:
            hidden_states = layer_module(hidden_states)
        return self.quantize_activations(hidden_states)


class QuantizedViTModel(QuantizedActivation):
    def __init__(self, orig, **quant_params):
        super().__init__(**quant_params)
        self.embeddings = QuantizedVitEmbeddings(orig.embeddings, **quant_params)
        self.encoder = QuantizedViTEncoder(orig.encoder, **quant_params)
        self.layernorm = _q(orig.layernorm, **quant_params)

    def forward(self, x):
        return self.layernorm(self.encoder(self.embeddings(x)))


class QuantizedVisionTransformerForImageClassification(QuantizedModel):
    def __init__(self, model_fp, input_size=(1, 3, 224, 224), quant_setup=None, **quant_params):
        super().__init__(input_size)
        self.vit = QuantizedViTModel(model_fp.vit, **quant_params)
        self.classifier = _q(model_fp.classifier, **quant_params)

    def forward(self, x):
        return self.classifier(self.vit(x)[:, 0, :])


def vit_quantized_approx(fp_model=None, **qparams):
    """models/vit_quantized_approx.py:396-398 with a seeded random-init ViT-B/16 (no checkpoint here)."""
    return QuantizedVisionTransformerForImageClassification(
        fp_model if fp_model is not None else VisionTransformerForImageClassification(), **qparams)
