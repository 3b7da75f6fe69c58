"""ORACLE / baseline support (test infrastructure, not product code): BASELINE config 1, leg C2 -- the reference's
`edgeEEGNetQuant` after `tq.convert`, on PyTorch's qnnpack engine, batch 1 on the host CPU.

/root/reference does not exist on the GPU box, so the module is RESTATED here (layer for layer, citing the reference):
  edge-eegnet_torch/edgeEEGnet_quant.py:7          engine 'qnnpack'
  edge-eegnet_torch/edgeEEGnet_quant.py:50-79      layers: QuantStub, ConvBnReLU2d(1, F1, (C, 1)), ConvBn2d(F1, F1, (1, 64),
                                                   padding (0, 31), groups F1), AvgPool (1, 8), Conv2d(F1, F1, (1, 16), padding
                                                   (0, 7), groups F1, no bias), ConvBnReLU2d(F1, F2, (1, 1)), AvgPool (1, 8),
                                                   Flatten, Linear(F2 * n_features, N), DeQuantStub;
                                                   n_features = (((T - 1) // 8) - 1) // 8
  edge-eegnet_torch/edgeEEGnet_quant.py:81-107     forward (dropout layers are identities at inference)
  edge-eegnet_torch/edgeEEGnet_controller.py:184-194, 216-217   qconfig (MinMaxObserver quint8 activations,
                                                   MinMaxObserver qint8 weights), prepare_qat, convert
The arithmetic itself lives in PyTorch's quantized backend (third party, version not pinned by the reference):
this file is a TIMING baseline, parity with it is unpinned and nothing compares logits against it.
"""
import time

import torch as t
import torch.ao.nn.intrinsic.qat as nniqat
import torch.ao.nn.qat as nnqat
import torch.ao.quantization as tq


class EdgeEEGNetQuantRestated(t.nn.Module):
    def __init__(self, qconfig, F1=16, D=1, F2=None, C=22, T=1125, N=4):
        super().__init__()
        if F2 is None:
            F2 = F1 * D
        self.qconfig = qconfig
        n_features = (((T - 1) // 8) - 1) // 8
        self.quant = tq.QuantStub(qconfig)
        self.conv1_bn = nniqat.ConvBnReLU2d(1, D * F1, (C, 1), momentum=0.01, qconfig=qconfig)
        self.conv2_bn_relu = nniqat.ConvBn2d(D * F1, D * F1, (1, 64), padding=(0, 31), groups=D * F1, momentum=0.01,
                                             qconfig=qconfig)
        self.pool1 = t.nn.AvgPool2d((1, 8))
        self.sep_conv1 = nnqat.Conv2d(D * F1, D * F1, (1, 16), groups=D * F1, bias=False, padding=(0, 7), qconfig=qconfig)
        self.sep_conv2_bn_relu = nniqat.ConvBnReLU2d(D * F1, F2, (1, 1), momentum=0.01, qconfig=qconfig)
        self.pool2 = t.nn.AvgPool2d((1, 8))
        self.flatten = t.nn.Flatten()
        self.fc = nnqat.Linear(F2 * n_features, N, bias=True, qconfig=qconfig)
        self.dequant = tq.DeQuantStub()

    def forward(self, x):
        x = x.reshape(x.shape[0], 1, x.shape[1], x.shape[2])
        x = self.quant(x)
        x = self.pool1(self.conv2_bn_relu(self.conv1_bn(x)))
        x = self.pool2(self.sep_conv2_bn_relu(self.sep_conv1(x)))
        return self.dequant(self.fc(self.flatten(x)))


def build_converted(C=64, T=480, seed=0):
    """Random-init module -> prepare_qat -> a few calibration passes -> convert (edgeEEGnet_controller.py:184-217)."""
    t.backends.quantized.engine = "qnnpack"
    g = t.Generator().manual_seed(seed)
    qconfig = tq.QConfig(activation=tq.MinMaxObserver.with_args(dtype=t.quint8), weight=tq.MinMaxObserver.with_args(dtype=t.qint8))
    m = EdgeEEGNetQuantRestated(qconfig, C=C, T=T)
    for p in m.parameters():
        if p.dim() > 1:
            p.data.copy_(t.empty_like(p).uniform_(-0.5, 0.5, generator=g))
    m.train()
    tq.prepare_qat(m, inplace=True)
    for _ in range(4):
        m(t.randn(8, C, T, generator=g))
    m.eval()
    tq.convert(m, inplace=True)
    return m


def time_batch1(C=64, T=480, threads=1, seconds=2.0, seed=0):
    """-> ms per trial at batch 1 (median-free mean over a bounded run), with `threads` intra-op threads."""
    m = build_converted(C, T, seed)
    prev = t.get_num_threads()
    if threads:
        t.set_num_threads(threads)
    x = t.randn(1, C, T, generator=t.Generator().manual_seed(seed + 1))
    with t.no_grad():
        for _ in range(5):
            y = m(x)
        n, t0 = 0, time.perf_counter()
        while time.perf_counter() - t0 < seconds:
            m(x)
            n += 1
        dt = time.perf_counter() - t0
    t.set_num_threads(prev)
    assert tuple(y.shape) == (1, 4)
    return 1e3 * dt / n, n


if __name__ == "__main__":
    for th in (1, 0):
        ms, n = time_batch1(threads=th)
        print(f"edgeEEGNetQuant (qnnpack, restated) 64x480 batch 1, threads={'default' if not th else th}: {ms:.3f} ms/trial over {n} calls")
