"""BERT-base masked-LM for BASELINE.json config 4 ("BERT-base MLM bf16 synthetic seq 128 (large gradient
buckets, fused all-reduce + update)"): 12 layers, hidden 768, 12 heads, FFN 3072, vocab 30522,
~110 M parameters, MLM decoder tied to the word embeddings.

Execution: TorchEngine. Every Dense projection (QKV, attention output, FFN, MLM transform - 73 of the
~110 M parameters... and all of the GEMM FLOPs except the tied decoder) runs forward, dgrad and wgrad on
the persistent TMA + tcgen05 GEMM under the mixed_bfloat16 policy (ops/gemm.py:linear); attention is the fused
tcgen05 kernel of ops/attention.py (at S = 128 a whole sequence fits one tile, so no sequence parallelism is
needed - SURVEY 5.7; other sequence lengths fall back to torch's scaled_dot_product_attention); the ~440 MB fp32 gradient arena is reduced by the
fused NVLink kernels. input = token ids [B, S] int, labels = token ids with -100 on unmasked positions.
"""
from __future__ import annotations

import math
import os
from typing import List, Sequence

import torch
import torch.nn.functional as F

from .. import keras
from ..keras import layers as L
from ..keras.layers import CallContext, Layer, truncated_normal


class BertMLM(keras.Model):
    def __init__(self, vocab_size: int = 30522, hidden: int = 768, layers: int = 12, heads: int = 12,
                 intermediate: int = 3072, max_position: int = 512, seq_len: int = 128, name: str = "bert") -> None:
        super().__init__(name)
        self.vocab_size, self.hidden, self.heads, self.seq_len = vocab_size, hidden, heads, seq_len
        init = truncated_normal(0.02)
        self._order: List[Layer] = []

        def add(l: Layer) -> Layer:
            self._order.append(l)
            return l

        self.word = add(L.Embedding(vocab_size, hidden, name="embeddings/word_embeddings"))
        self.pos = add(L.Embedding(max_position, hidden, name="embeddings/position_embeddings"))
        self.typ = add(L.Embedding(2, hidden, name="embeddings/token_type_embeddings"))
        self.emb_ln = add(L.LayerNormalization(name="embeddings/layer_norm"))
        self.blocks = []
        for i in range(layers):
            p = f"encoder/layer_{i}"
            self.blocks.append(dict(
                q=add(L.Dense(hidden, kernel_initializer=init, name=f"{p}/attention/query")),
                k=add(L.Dense(hidden, kernel_initializer=init, name=f"{p}/attention/key")),
                v=add(L.Dense(hidden, kernel_initializer=init, name=f"{p}/attention/value")),
                o=add(L.Dense(hidden, kernel_initializer=init, name=f"{p}/attention/output")),
                ln1=add(L.LayerNormalization(name=f"{p}/attention/layer_norm")),
                ff1=add(L.Dense(intermediate, activation="gelu", kernel_initializer=init, name=f"{p}/intermediate")),
                ff2=add(L.Dense(hidden, kernel_initializer=init, name=f"{p}/output")),
                ln2=add(L.LayerNormalization(name=f"{p}/output_layer_norm")),
            ))
        self.mlm_dense = add(L.Dense(hidden, activation="gelu", kernel_initializer=init, name="mlm/transform"))
        self.mlm_ln = add(L.LayerNormalization(name="mlm/layer_norm"))
        self.mlm_bias = add(_Bias(vocab_size, name="mlm/output_bias"))
        self.build((seq_len,))

    def _layers(self) -> List[Layer]:
        return self._order

    def build(self, input_shape: Sequence[int]) -> None:
        if self.built:
            return
        S, H = int(input_shape[0]), self.hidden
        self._input_shape = (S,)
        for l in (self.word, self.pos, self.typ):
            l._maybe_build((S,))
        self.emb_ln._maybe_build((S, H))
        for b in self.blocks:
            for k in ("q", "k", "v", "o", "ff1"):
                b[k]._maybe_build((S, H))
            b["ff2"]._maybe_build((S, b["ff1"].units))
            b["ln1"]._maybe_build((S, H))
            b["ln2"]._maybe_build((S, H))
        self.mlm_dense._maybe_build((S, H))
        self.mlm_ln._maybe_build((S, H))
        self.mlm_bias._maybe_build((S, self.vocab_size))
        self.built = True

    def _forward(self, ids: torch.Tensor, ctx: CallContext) -> torch.Tensor:
        B, S = ids.shape
        dt = ctx.compute_dtype
        pos = torch.arange(S, device=ids.device)
        x = self.word.call(ids, ctx) + self.pos.call(pos, ctx)[None] + ctx.param(self.typ.embeddings)[0]
        x = self.emb_ln.call(x, ctx).to(dt)
        nh, hd = self.heads, self.hidden // self.heads
        fuse_qkv = bool(x.is_cuda and dt == torch.bfloat16 and self.hidden % 8 == 0 and (B * S) % 8 == 0)
        native_attn = False
        if fuse_qkv:
            from ..ops.attention import attention as _attention
            from ..ops.gemm import linear as _tc_linear

            native_attn = S == 128 and hd == 64 and os.environ.get("TFDL_ATTENTION", "native") == "native"
        for b in self.blocks:
            if fuse_qkv:
                # query / key / value stay three Keras variables (names, checkpoints); for the tensor core they are one
                # [3H, H] operand: a single N = 3H GEMM fills the machine where three N = H ones leave 1.3 waves each
                w = torch.cat([ctx.param(b[n].kernel) for n in ("q", "k", "v")], 0)
                bias = torch.cat([ctx.param(b[n].bias) for n in ("q", "k", "v")], 0)
                qkv = _tc_linear(x, w, bias, None)
                if native_attn:
                    # per-head Q / K / V tiles are cut out of the fused projection by TMA inside the kernel
                    a = _attention(qkv, nh)
                    x = b["ln1"].call(b["o"].call(a, ctx), ctx, residual=x).to(dt)
                    x = b["ln2"].call(b["ff2"].call(b["ff1"].call(x, ctx), ctx), ctx, residual=x).to(dt)
                    continue
                q, k, v = (t.view(B, S, nh, hd).transpose(1, 2) for t in qkv.split(self.hidden, -1))
            else:
                q = b["q"].call(x, ctx).view(B, S, nh, hd).transpose(1, 2)
                k = b["k"].call(x, ctx).view(B, S, nh, hd).transpose(1, 2)
                v = b["v"].call(x, ctx).view(B, S, nh, hd).transpose(1, 2)
            a = F.scaled_dot_product_attention(q, k, v).transpose(1, 2).reshape(B, S, self.hidden)
            x = b["ln1"].call(b["o"].call(a, ctx), ctx, residual=x).to(dt)
            x = b["ln2"].call(b["ff2"].call(b["ff1"].call(x, ctx), ctx), ctx, residual=x).to(dt)
        h = self.mlm_ln.call(self.mlm_dense.call(x, ctx), ctx).to(dt)
        w = ctx.param(self.word.embeddings)  # tied decoder
        if dt == torch.bfloat16 and h.is_cuda:
            from ..ops.gemm import tied_decoder

            # bf16 logits (a column slice of the padded GEMM output): the loss kernel reads them as they are
            return tied_decoder(h, w, ctx.param(self.mlm_bias.bias))
        logits = F.linear(h, w.to(dt)) if dt != torch.float32 else F.linear(h, w)
        return self.mlm_bias.call(logits.float(), ctx)


class _Bias(Layer):
    def __init__(self, n: int, **kw) -> None:
        super().__init__(**kw)
        self.n = n

    def build(self, input_shape) -> None:
        self.bias = self.add_weight("bias", (self.n,), "zeros")

    def call(self, x: torch.Tensor, ctx: CallContext) -> torch.Tensor:
        return x + ctx.param(self.bias)


def build_and_compile_bert_mlm(learning_rate: float = 1e-4, seq_len: int = 128, optimizer: str = "sgd", **kw) -> BertMLM:
    model = BertMLM(seq_len=seq_len, **kw)
    opt = keras.optimizers.Adam(learning_rate) if optimizer == "adam" else keras.optimizers.SGD(learning_rate, momentum=0.9)
    model.compile(loss=keras.losses.SparseCategoricalCrossentropy(from_logits=True, ignore_class=-100), optimizer=opt)
    return model
