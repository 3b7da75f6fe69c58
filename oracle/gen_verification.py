#!/usr/bin/env python
"""Generates tests/golden/export_<tag>/{net.npz, config.json, verification.npz} by running the REFERENCE's QuantLab
`edgeEEGNet` module (/root/reference/QuantLab/quantlab/PhysionetMMMI/edgeEEGNet/edgeEEGnet.py:137-168, float-simulated
quantised forward, eval mode) on a synthetic state_dict, i.e. the three files `QuantLab/export_net_data.py:85-87,112-170`
leaves in an export folder.  Test infrastructure; runs only in the build container (no reference tree on the GPU box).

The tap names of verification.npz are those of export_net_data.py:112-167; here every tap is the output of the module of
the same role, captured with forward hooks (module -> tap name in TAP_OF_MODULE).  The reference's own consumer of this
file is edge-eegnet_wolf/test/python/golden_model/testcase.py:67-93 (integer model within TOLERANCE quantisation steps
of the float-simulated taps) -- mibminet_b200.verify mirrors that check on the GPU path."""
import json
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)
REF = "/root/reference/QuantLab/quantlab"

TAP_OF_MODULE = {
    "quant1": "input_quant",
    "conv1": "layer1_conv_out", "batch_norm1": "layer1_bn_out", "quant2": "layer1_activ",
    "conv2": "layer2_conv_out", "batch_norm2": "layer2_bn_out", "activation1": "layer2_relu_out",
    "pool1": "layer2_pool_out", "quant3": "layer2_activ",
    "sep_conv1": "layer3_conv_out", "quant4": "layer3_activ",
    "sep_conv2": "layer4_conv_out", "batch_norm3": "layer4_bn_out", "activation2": "layer4_relu_out",
    "pool2": "layer4_pool_out", "quant5": "layer4_activ",
    "fc": "output", "quant6": "output_quant",
}

# (tag, C, T, seed, module file)
CASES = (("c64_t480", 64, 480, 21, f"{REF}/PhysionetMMMI/edgeEEGNet/edgeEEGnet.py"),
         ("c22_t1125", 22, 1125, 22, f"{REF}/BCI-CompIV-2a/edgeEEGNet/edgeEEGnet.py"))


def load_reference_edge_eegnet(path):
    import importlib.util
    from gen_golden_eegnet import load_reference_eegnet
    load_reference_eegnet()                      # injects quantlab.indiv (+ inq_ops / ste_ops), SURVEY 8c O3
    spec = importlib.util.spec_from_file_location("ref_edge_eegnet_" + str(abs(hash(path))), path)
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m.edgeEEGNet


def main():
    from mibminet_b200 import synth
    for tag, C, T, seed, path in CASES:
        edgeEEGNet = load_reference_edge_eegnet(path)
        sd = synth.make_state_dict(C=C, T=T, seed=seed, negative_gamma=False)
        cfg = synth.make_config(C=C, T=T)
        net = edgeEEGNet(F1=16, D=1, C=C, T=T, N=4, Nf=64, weightInqSchedule={"1": 1.0}, floorToZero=True,
                         actSTEStartEpoch=0)
        res = net.load_state_dict({k: torch.from_numpy(np.asarray(v)) for k, v in sd.items()}, strict=False)
        assert not res.unexpected_keys, res
        net.eval()
        for m in net.modules():                  # activations quantise, frozen (INQ-grid) weights are the weights
            if hasattr(m, "started"):
                m.started, m.monitoring = True, False
            if hasattr(m, "weightFrozen") and hasattr(m, "weight"):
                m.weight.data.copy_(m.weightFrozen.data)

        taps, hooks = {}, []
        for name, tap in TAP_OF_MODULE.items():
            hooks.append(getattr(net, name).register_forward_hook(
                lambda _m, _i, out, tap=tap: taps.__setitem__(tap, out.detach().numpy().copy())))
        rng = np.random.default_rng(seed)
        s1 = float(sd["quant1.absMaxValue"][0])
        x = (rng.standard_normal((C, T)) * 0.35 * s1).astype(np.float32)      # one float trial, a few values clip
        with torch.no_grad():
            net(torch.from_numpy(x).reshape(1, 1, C, T))
        for h in hooks:
            h.remove()
        taps["input"] = x

        d = os.path.join(ROOT, "tests", "golden", "export_" + tag)
        os.makedirs(d, exist_ok=True)
        full = {k: v.detach().numpy() for k, v in net.state_dict().items()}   # every key the checkpoint would hold
        np.savez_compressed(os.path.join(d, "net.npz"), **full)
        with open(os.path.join(d, "config.json"), "w") as f:
            json.dump(cfg, f, indent=1)
        np.savez_compressed(os.path.join(d, "verification.npz"), **{k: v.astype(np.float32) for k, v in taps.items()})
        print(tag, "taps:", sorted(taps), "missing keys at load:", res.missing_keys)


if __name__ == "__main__":
    main()
