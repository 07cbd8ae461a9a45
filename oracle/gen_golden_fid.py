"""Generates tests/golden/fid_resnet_ref.npz from the UNMODIFIED reference FID_ResNet.py
(imports cleanly, no stubs).   PYTHONDONTWRITEBYTECODE=1 python oracle/gen_golden_fid.py"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "tests", "golden", "fid_resnet_ref.npz")
sys.dont_write_bytecode = True
sys.path.insert(0, "/root/reference")
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import FID_ResNet as ref          # noqa: E402  (the reference)
import ref_fid_resnet as port     # noqa: E402

torch.manual_seed(31)
m = ref.resnet50()
torch.manual_seed(31)
p = port.resnet50()
sd_r, sd_p = m.state_dict(), p.state_dict()
assert list(sd_r) == list(sd_p), "state_dict keys / order differ"
assert all(torch.equal(sd_r[k], sd_p[k]) for k in sd_r), "seed replay does not reproduce the reference weights"
torch.manual_seed(32)
x = torch.randn(2, 1, 32, 32, 32)
with torch.no_grad():                       # BN stays in train mode
    y = m(x)
feat = y.mean(dim=(2, 3, 4))                # eval_utils.py:43-44
np.savez_compressed(OUT, torch_version=np.array(torch.__version__), x_seed=32, w_seed=31,
                    y_shape=np.array(y.shape), y_sample=y[:, ::64].numpy(), feat=feat.numpy(),
                    conv1_w_sum=sd_r["conv1.weight"].double().sum().item(),
                    l4_w_sum=sd_r["layer4.2.conv3.weight"].double().sum().item())
print("wrote", OUT, os.path.getsize(OUT), "y", tuple(y.shape), "feat[0,:4]", feat[0, :4])
