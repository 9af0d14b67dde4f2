"""TEST / BASELINE INFRASTRUCTURE ONLY -- the reference model's forward, executed by torch on the CPU.

/root/reference/src/modules/model.py (Model.forward :137-156, residual_layer :93-105, rnn_layer :124-135) restated as a
function over ANY module that carries the reference's attribute names (batchNorm, incpConv0/1, conv0..3, cnn_fc_rnn, the
unregistered `rnns` list of (name, nn.GRU), fc1) -- deepore_b200.model.Model holds exactly these nn modules, so
`build(state)` gives the nn.Conv2d / nn.BatchNorm2d / nn.Linear / nn.GRU stack the reference runs, filled from a numpy
checkpoint.  This is the schedule and the arithmetic (torch CPU fp32, oneDNN / MKL kernels) of the reference's own
`test.py` run; bench.py times it as the `torch_cpu` leg of the CPU baseline, tests/test_model_oracle.py pins it on the
golden logits of the unmodified reference class.  The reference cannot travel to the GPU box (it needs /root/reference),
this restatement can.  Never imported by the product package.
"""
from __future__ import annotations

import numpy as np
import torch


def build(state, in_ch=4, width=200):
    """nn modules of the reference architecture (parity configuration) filled from a numpy state, eval mode, CPU."""
    import os
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    if root not in sys.path:
        sys.path.insert(0, root)
    from deepore_b200 import model as dmodel, modelio
    m = dmodel.Model(input_channel=in_ch, coverage_depth=width)
    modelio.load_into_module(m, state)
    m.eval()
    return m


def residual_layer(m, x, layer, batch_norm_flag=False):            # model.py:93-105
    incp = m.incpConv1 if layer != 0 else m.incpConv0
    conv1 = m.conv1 if layer != 0 else m.conv0
    skip = m.batchNorm(incp(x)) if batch_norm_flag else incp(x)
    c1 = m.batchNorm(conv1(x))
    c2 = m.batchNorm(m.conv2(c1))
    return skip + m.conv3(c2)


def forward(m, x, hidden):
    """x (N, C, 3, W) float32 = ToTensor of three image rows, hidden (2, N, H) zeros -> logits (N, 3, class_n)."""
    x = residual_layer(m, x, 0, True)                              # model.py:138-142
    for layer in range(1, 5):
        x = residual_layer(m, x, layer)
    n, c, t, w = x.size()
    x = x.view(n, c, w, t).view(n, c * w, t)                       # model.py:144-149: reinterpreting views, no transpose
    x = x.transpose(1, 2).transpose(0, 1).contiguous()             # T x N x (C*W)
    x = m.cnn_fc_rnn(x)                                            # model.py:151
    for _, rnn in m.rnns:                                          # model.py:124-135: hidden of layer k feeds layer k+1
        x, hidden = rnn(x, hidden)
        x = x.view(x.size(0), x.size(1), 2, -1).sum(2).view(x.size(0), x.size(1), -1)
    return m.fc1(x).transpose(0, 1)                                # model.py:154-155


def forward_image_rows(m, images_u8, rows, batch=None):
    """The reference's schedule (test.py:59-66, naive: the CNN runs on all three rows of every call): images_u8
    [N, H, W, C] uint8, `rows` = first image rows of the calls to run -> logits [N, len(rows), 3, class_n]."""
    images = (torch.from_numpy(np.ascontiguousarray(images_u8)).permute(0, 3, 1, 2).float() / 255.0).contiguous()
    n = images.size(0)
    hidden = torch.zeros(2, n, m.hidden_size)
    out = []
    with torch.no_grad():
        for r in rows:
            out.append(forward(m, images[:, :, r:r + 3, :], hidden))
    return torch.stack(out, dim=1).numpy()
