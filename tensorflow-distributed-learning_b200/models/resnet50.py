"""Keras-applications ResNet-50 (v1) for BASELINE.json config 3 ("Keras ResNet-50 bf16 synthetic 224x224x3,
batch 256/GPU"). Same topology, layer names and parameter count as tf.keras.applications.ResNet50
(25,636,712 parameters, 53,120 of them non-trainable BatchNorm statistics), NHWC.

Execution: this model runs on the TorchEngine with every layer op on this library's kernels: the 3x3 / strided / 7x7
convolutions on the tcgen05 implicit-GEMM kernel (TMA im2col loads; the 3-channel stem through a patch matrix + GEMM),
1x1 convolutions and the classifier on the TMA + tcgen05 GEMM, max-pool and global average pool on streaming kernels,
BatchNorm + ReLU + the residual add fused in the kernels of ops/batchnorm.py, and the ~102 MB gradient arena reduced in
buckets by the fused NVLink all-reduce + SGD/momentum kernels while backward is still running (ops/allreduce.py).
"""
from __future__ import annotations

from typing import List, Optional, Sequence, Tuple

import torch

from .. import keras
from ..keras import layers as L
from ..keras.layers import CallContext, Layer


class ResNet50(keras.Model):
    def __init__(self, classes: int = 1000, input_shape: Sequence[int] = (224, 224, 3), name: str = "resnet50") -> None:
        super().__init__(name)
        self.classes = classes
        self._order: List[Layer] = []
        self._blocks: List[Tuple[str, List[Layer], Optional[List[Layer]]]] = []
        he = "he_normal"

        def conv(name, f, k, s=1, pad="valid"):
            l = L.Conv2D(f, k, strides=s, padding=pad, name=name, kernel_initializer=he)
            self._order.append(l)
            return l

        def bn(name):
            l = L.BatchNormalization(epsilon=1.001e-5, name=name)
            self._order.append(l)
            return l

        self.stem = [conv("conv1_conv", 64, 7, 2, "same"), bn("conv1_bn")]
        self.pool1 = L.MaxPooling2D(3, 2, padding="same", name="pool1_pool")
        self._order.append(self.pool1)
        for stage, (filters, blocks, stride) in enumerate([(64, 3, 1), (128, 4, 2), (256, 6, 2), (512, 3, 2)], start=2):
            for b in range(1, blocks + 1):
                p = f"conv{stage}_block{b}"
                s = stride if b == 1 else 1
                shortcut = [conv(f"{p}_0_conv", 4 * filters, 1, s), bn(f"{p}_0_bn")] if b == 1 else None
                main = [conv(f"{p}_1_conv", filters, 1, s), bn(f"{p}_1_bn"),
                        conv(f"{p}_2_conv", filters, 3, 1, "same"), bn(f"{p}_2_bn"),
                        conv(f"{p}_3_conv", 4 * filters, 1), bn(f"{p}_3_bn")]
                self._blocks.append((p, main, shortcut))
        self.avg = L.GlobalAveragePooling2D(name="avg_pool")
        self.fc = L.Dense(classes, name="predictions")
        self._order += [self.avg, self.fc]
        self.build(tuple(input_shape))

    def _layers(self) -> List[Layer]:
        return self._order

    def build(self, input_shape: Sequence[int]) -> None:
        if self.built:
            return
        shape = tuple(int(s) for s in input_shape)
        self._input_shape = shape
        for l in self.stem:
            shape = l._maybe_build(shape)
        shape = self.pool1._maybe_build(shape)
        for _, main, shortcut in self._blocks:
            s_in = shape
            if shortcut is not None:
                ss = s_in
                for l in shortcut:
                    ss = l._maybe_build(ss)
            for l in main:
                shape = l._maybe_build(shape)
        shape = self.avg._maybe_build(shape)
        self.fc._maybe_build(shape)
        self.built = True

    def _forward(self, x: torch.Tensor, ctx: CallContext) -> torch.Tensor:
        if ctx.compute_dtype != torch.float32:
            x = x.to(ctx.compute_dtype)
        # Conv bias, BatchNorm, ReLU and the bottleneck's residual add are ONE fused pass over the activations: the
        # convolution leaves its bias out and the BatchNorm kernel absorbs it (keras/layers.py:BatchNormalization)
        def cbn(conv: Layer, bn: Layer, t: torch.Tensor, relu: bool, residual: Optional[torch.Tensor] = None) -> torch.Tensor:
            bias = ctx.param(conv.bias) if conv.bias is not None else None
            return bn.call(conv.call(t, ctx, add_bias=False), ctx, relu=relu, residual=residual, in_bias=bias)

        x = cbn(self.stem[0], self.stem[1], x, True)
        x = self.pool1.call(x, ctx)
        for _, main, shortcut in self._blocks:
            sc = x if shortcut is None else cbn(shortcut[0], shortcut[1], x, False)
            y = cbn(main[0], main[1], x, True)
            y = cbn(main[2], main[3], y, True)
            x = cbn(main[4], main[5], y, True, sc)
        return self.fc.call(self.avg.call(x, ctx), ctx).float()


def build_and_compile_resnet50(learning_rate: float = 0.1, momentum: float = 0.9, classes: int = 1000,
                               input_shape: Sequence[int] = (224, 224, 3)) -> ResNet50:
    model = ResNet50(classes=classes, input_shape=input_shape)
    model.compile(loss=keras.losses.SparseCategoricalCrossentropy(from_logits=True),
                  optimizer=keras.optimizers.SGD(learning_rate=learning_rate, momentum=momentum),
                  metrics=[keras.metrics.SparseCategoricalAccuracy()])
    return model
