#!/usr/bin/env python
"""Generates tests/golden/eegnet_order_*.npz by running the REFERENCE's QuantLab EEGNet module
(/root/reference/QuantLab/quantlab/PhysionetMMMI/EEGNet/eegnet.py, float-simulated quantised forward, eval mode) on the
synthetic state_dict of mibminet_b200.synth.make_eegnet_state_dict.  Test infrastructure; runs only in the build
container (the reference tree does not exist on the GPU box).  The fixtures anchor oracle/eegnet_np.float_sim (and
through it the integer pipeline, within the reference's 10-LSB tolerance) to the reference's own code.

Recipe: the tree ships no __init__.py files, so `quantlab.indiv` is injected with the `Controller` base
class the ops import, and the module files are loaded by path."""
import importlib.util
import os
import sys
import types

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
REF = "/root/reference/QuantLab/quantlab"


def load_reference_eegnet():
    class Controller(object):
        def __init__(self):
            pass

        def step_preTraining(self, *a, **k):
            pass

        def step_preValidation(self, *a, **k):
            pass

        def step_postOptimStep(self, *a, **k):
            pass

        @staticmethod
        def getControllers(net):
            return [v for v in net.__dict__.values() if isinstance(v, Controller)]

    ql, indiv = types.ModuleType("quantlab"), types.ModuleType("quantlab.indiv")
    indiv.Controller = Controller
    ql.indiv = indiv
    sys.modules["quantlab"], sys.modules["quantlab.indiv"] = ql, indiv

    def load(name, path):
        spec = importlib.util.spec_from_file_location(name, path)
        m = importlib.util.module_from_spec(spec)
        sys.modules[name] = m
        spec.loader.exec_module(m)
        return m

    indiv.inq_ops = load("quantlab.indiv.inq_ops", f"{REF}/indiv/inq_ops.py")
    indiv.ste_ops = load("quantlab.indiv.ste_ops", f"{REF}/indiv/ste_ops.py")
    return load("ref_eegnet", f"{REF}/PhysionetMMMI/EEGNet/eegnet.py").EEGNet


def main():
    from mibminet_b200 import synth
    EEGNet = load_reference_eegnet()
    out_dir = os.path.join(ROOT, "tests", "golden")
    for tag, C, T, F1, D, B in (("c22_t1125", 22, 1125, 8, 2, 3), ("c64_t480", 64, 480, 8, 2, 3)):
        sd = synth.make_eegnet_state_dict(C=C, T=T, F1=F1, D=D, seed=11, negative_gamma=False)
        net = EEGNet(F1=F1, D=D, C=C, T=T, N=4, weightInqSchedule={"1": 1.0}, floorToZero=True, actSTEStartEpoch=0)
        missing = net.load_state_dict({k: torch.from_numpy(np.asarray(v)) for k, v in sd.items()}, strict=False)
        assert not missing.unexpected_keys, missing
        net.eval()
        for m in net.modules():                      # activations quantise (started), frozen weights are used
            if hasattr(m, "started"):
                m.started, m.monitoring = True, False
            if hasattr(m, "weightFrozen") and hasattr(m, "weight"):
                m.weight.data.copy_(m.weightFrozen.data)     # all weights frozen on the INQ grid = the exported net
        x_int = synth.make_trials(B, C, T, seed=5)
        x = torch.from_numpy(x_int.astype(np.float32) / 127.0 * float(sd["quant1.absMaxValue"][0]))[:, None]
        with torch.no_grad():
            y = net(x).numpy()
        s6 = float(sd["quant6.absMaxValue"][0])
        np.savez_compressed(os.path.join(out_dir, f"eegnet_order_{tag}.npz"), x_int=x_int, logits_float=y,
                            logits_units=y / s6 * 127.0, C=C, T=T, F1=F1, D=D, seed=11, missing=np.array(missing.missing_keys))
        print(tag, "reference logits (int units):", np.round(y / s6 * 127.0, 2)[0], "missing keys:", missing.missing_keys)


if __name__ == "__main__":
    main()
