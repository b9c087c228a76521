"""CPU restatement of the reference's BADGESelector gradient embedding (oracle side; TEST INFRASTRUCTURE ONLY).

Path restated (reference file:line):
  * `BADGESelector.cal_scores` (src/activelearning/badge_selector.py:56-104): per pool image (batch_size is forced to 1,
    al_trainer.py:875) `outputs = model(image)`, pseudo-labels `preds = outputs.softmax(1).argmax(1)` (:79-80),
    `loss = ce_loss(outputs, preds) + dice_loss(outputs, preds)` (:91-95, the only reachable branch,
    `multiple_loss="add"`, al_trainer.py:878), embedding = gradient of that loss w.r.t.
    `decoder.seg_output.weight` flattened to [K*32] (`image_wise_grad`, :19-34).
  * losses as `al_train` configures them (al_trainer.py:782-798): `DiceLoss(num_classes, smooth=1e-5, do_bg=True,
    softmax=True, batch=False, squared=False)` (src/losses/dice_loss.py:7-76: per class
    `1 - (2*sum(p*y) + s) / (sum(p) + sum(y) + s)`, mean over classes) and `torch.nn.CrossEntropyLoss()` (mean over
    pixels).
  * `select_next_batch` (:106-128): seeded-random cold start when nothing is labeled, else
    `kmeans_plusplus(X=grad_embed, n_clusters=select_num)` (unweighted, numpy global RNG) -> names in pick order.

`grad_embedding_autograd` follows the reference literally (torch autograd through the 1x1 head only - the features
below it do not depend on the head weight); `grad_embedding_closed_form` is the formula the CUDA kernels evaluate:
    g_k(px) = (p_k - y_k)/P + p_k * (q_k - sum_j p_j q_j),   q_c = (-2*y_c/U_c + (2*I_c + s)/U_c^2) / C
    dL/dW[k, c] = sum_px g_k(px) * feat_c(px)
with I_c = sum p_c*y_c, U_c = sum p_c + sum y_c + s, C = number of dice classes.
"""
from __future__ import annotations

import numpy as np
import torch
import torch.nn.functional as F

from .unet_ref import unet_forward


def dice_loss(outputs: torch.Tensor, labels: torch.Tensor, smooth: float = 1e-5, do_bg: bool = True) -> torch.Tensor:
    """DiceLoss.forward (dice_loss.py:32-76) with softmax=True, batch=False, squared=False."""
    p = torch.softmax(outputs, dim=1)
    y = F.one_hot(labels.long(), outputs.shape[1]).permute(0, 3, 1, 2).float()
    if not do_bg:
        p, y = p[:, 1:], y[:, 1:]
    inter = (p * y).sum((2, 3))
    dice = 1 - (2 * inter + smooth) / (p.sum((2, 3)) + y.sum((2, 3)) + smooth)
    return dice.mean()


def grad_embedding_autograd(sd, x: torch.Tensor, smooth: float = 1e-5, do_bg: bool = True) -> np.ndarray:
    """[n, K*32] embeddings, one image at a time like the reference (x: [n,1,H,W] fp32)."""
    out = []
    for i in range(x.shape[0]):
        with torch.no_grad():
            _, feat, _ = unet_forward(sd, x[i:i + 1], return_features=True)
        w = sd["decoder.seg_output.weight"].detach().clone().requires_grad_(True)
        logits = F.conv2d(feat, w, sd["decoder.seg_output.bias"])
        preds = logits.softmax(1).argmax(1)
        loss = F.cross_entropy(logits, preds) + dice_loss(logits, preds, smooth, do_bg)
        (g,) = torch.autograd.grad(loss, w)
        out.append(g.flatten().numpy())
    return np.stack(out).astype(np.float32)


def grad_embedding_closed_form(sd, x: torch.Tensor, smooth: float = 1e-5, do_bg: bool = True) -> np.ndarray:
    """Same embedding from the closed form (fp64), the formula of csrc/badge.cuh."""
    out = []
    with torch.no_grad():
        for i in range(x.shape[0]):
            logits, feat, _ = unet_forward(sd, x[i:i + 1], return_features=True)
            K = logits.shape[1]
            z = logits[0].reshape(K, -1).double()
            f = feat[0].reshape(feat.shape[1], -1).double()
            P = z.shape[1]
            p = torch.softmax(z, 0)
            y = F.one_hot(torch.softmax(logits, 1).argmax(1).reshape(-1), K).T.double()
            I, U = (p * y).sum(1), p.sum(1) + y.sum(1) + smooth
            C = K if do_bg else K - 1
            q = (-2 * y / U[:, None] + ((2 * I + smooth) / U ** 2)[:, None]) / C
            if not do_bg:
                q[0] = 0
            g = (p - y) / P + p * (q - (p * q).sum(0, keepdim=True))
            out.append((g @ f.T).flatten().numpy())
    return np.stack(out).astype(np.float32)
