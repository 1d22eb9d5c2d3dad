"""BERT-base caller of the NON-approx path (BASELINE.json configs[4]).

The reference's models/bert_quantized.py:9-28 quantizes nothing (it downloads bert-base-uncased, runs
one sentence and returns the FP32 model), so there is no model-level reference; as SURVEY.md
appendix A prescribes, the model is a BERT-base graph (`BertConfig()` defaults: 12 layers, hidden 768,
MLP 3072, 12 heads, vocab 30522) passed through `quantize_model` (approx/replace_operations_with_approx_ops.py
:345-384) with run_method = {approx_flag: False, quantize_after_mult_and_add: False,
res_quantizer_flag: True, original_quantize_res: True}: every nn.Linear becomes a QCustomLinearTorch
whose forward is  fake-quant(x) , fake-quant(W) , x @ W^T on the tcgen05 tensor cores
(csrc/fq_gemm_tc.cu) , fake-quant(result); every LayerNorm becomes QuantLayerNorm.  Parity is pinned at
operator level (tests/test_gpu_fq_gemm_tc.py).  Weights are seeded random (no network)."""
import torch
from torch import nn
import torch.nn.functional as F

from .model_wrap import QuantizedModel, quantize_model

RUN_PLAIN = dict(approx_flag=False, quantize_after_mult_and_add=False, res_quantizer_flag=True,
                 original_quantize_res=True)


class BertConfig:
    def __init__(self, vocab_size=30522, hidden_size=768, num_hidden_layers=12, num_attention_heads=12,
                 intermediate_size=3072, max_position_embeddings=512, type_vocab_size=2, layer_norm_eps=1e-12):
        self.__dict__.update(locals())
        del self.__dict__["self"]


class BertEmbeddings(nn.Module):
    def __init__(self, c):
        super().__init__()
        self.word_embeddings = nn.Embedding(c.vocab_size, c.hidden_size)
        self.position_embeddings = nn.Embedding(c.max_position_embeddings, c.hidden_size)
        self.token_type_embeddings = nn.Embedding(c.type_vocab_size, c.hidden_size)
        self.LayerNorm = nn.LayerNorm(c.hidden_size, eps=c.layer_norm_eps)

    def forward(self, input_ids):
        pos = torch.arange(input_ids.shape[1], device=input_ids.device)
        e = self.word_embeddings(input_ids) + self.position_embeddings(pos)[None] + self.token_type_embeddings.weight[0]
        return self.LayerNorm(e)


class BertSelfAttention(nn.Module):
    def __init__(self, c):
        super().__init__()
        self.heads, self.head_size = c.num_attention_heads, c.hidden_size // c.num_attention_heads
        self.query, self.key, self.value = (nn.Linear(c.hidden_size, c.hidden_size) for _ in range(3))

    def forward(self, x):
        split = lambda t: t.view(t.shape[:-1] + (self.heads, self.head_size)).permute(0, 2, 1, 3)
        ctx = F.scaled_dot_product_attention(split(self.query(x)), split(self.key(x)), split(self.value(x)))
        ctx = ctx.permute(0, 2, 1, 3).contiguous()
        return ctx.view(ctx.shape[:-2] + (self.heads * self.head_size,))


class BertLayer(nn.Module):
    def __init__(self, c):
        super().__init__()
        self.attention = BertSelfAttention(c)
        self.attention_output = nn.Linear(c.hidden_size, c.hidden_size)
        self.attention_norm = nn.LayerNorm(c.hidden_size, eps=c.layer_norm_eps)
        self.intermediate = nn.Linear(c.hidden_size, c.intermediate_size)
        self.output = nn.Linear(c.intermediate_size, c.hidden_size)
        self.output_norm = nn.LayerNorm(c.hidden_size, eps=c.layer_norm_eps)

    def forward(self, h):
        h = self.attention_norm(self.attention_output(self.attention(h)) + h)
        return self.output_norm(self.output(F.gelu(self.intermediate(h))) + h)


class BertModel(nn.Module):
    def __init__(self, config=None):
        super().__init__()
        self.config = c = config or BertConfig()
        self.embeddings = BertEmbeddings(c)
        self.layer = nn.ModuleList(BertLayer(c) for _ in range(c.num_hidden_layers))
        self.pooler = nn.Linear(c.hidden_size, c.hidden_size)
        for m in self.modules():                      # BERT init: N(0, 0.02), zero bias
            if isinstance(m, (nn.Linear, nn.Embedding)):
                m.weight.data.normal_(0.0, 0.02)
            if isinstance(m, nn.Linear):
                m.bias.data.zero_()

    def forward(self, input_ids):
        h = self.embeddings(input_ids)
        for layer in self.layer:
            h = layer(h)
        return torch.tanh(self.pooler(h[:, 0]))


class QuantizedBert(QuantizedModel):
    def __init__(self, model_fp, **quant_params):
        super().__init__(input_size=(1, 128))
        self.bert = quantize_model(model_fp, **quant_params)

    def forward(self, input_ids):
        return self.bert(input_ids)


def bert_quantized(fp_model=None, **qparams):
    qparams.setdefault("run_method", dict(RUN_PLAIN))
    return QuantizedBert(fp_model if fp_model is not None else BertModel(), **qparams)
