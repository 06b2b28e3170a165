"""Mask R-CNN ResNet-50-FPN on the b200flow detection kernels.

The reference trains tensorpack's and aws-samples' TensorFlow Mask R-CNN under Horovod
(examples/maskrcnn/train-maskrcnn-tensorpack.yaml:17-35, train-maskrcnn-aws.yaml:17-37;
graph in SURVEY.md §3.4): ResNet-50 backbone with frozen BN, FPN, RPN head, proposal
generation (decode + per-level top-k + NMS), fg/bg sampling, multilevel RoIAlign, 2-fc box
head, 4-conv + deconv mask head.  Same architecture here in PyTorch: convolutions stay on
cuDNN (channels_last, bf16) with the frozen batch norm folded into their weights
(models/resnet_fpn.py), everything the survey lists as a kernel —
RoIAlign (K12), box decode + batched NMS (K13), matcher / IoU (K14), mask crop / paste (K15),
the fc layers (K0) — runs on b200flow kernels.
"""
from __future__ import annotations

import math
from dataclasses import dataclass

import torch
import torch.nn.functional as F
from torch import nn

from ..ops import detection as D
from ..ops import gemm as G


@dataclass
class MaskRCNNConfig:
    num_classes: int = 81                 # COCO 80 + background
    mode_mask: bool = True
    mode_fpn: bool = True
    backbone: str = "resnet50"
    anchor_sizes: tuple = (32, 64, 128, 256, 512)
    anchor_ratios: tuple = (0.5, 1.0, 2.0)
    rpn_pre_nms_train: int = 2000
    rpn_pre_nms_test: int = 1000
    rpn_post_nms_train: int = 2000
    rpn_post_nms_test: int = 1000
    rpn_nms_thresh: float = 0.7
    rpn_fg_iou: float = 0.7
    rpn_bg_iou: float = 0.3
    rpn_batch: int = 256
    roi_batch: int = 512
    roi_fg_fraction: float = 0.25
    roi_fg_iou: float = 0.5
    box_weights: tuple = (10.0, 10.0, 5.0, 5.0)
    det_score_thresh: float = 0.05
    det_nms_thresh: float = 0.5
    det_per_image: int = 100
    fc_dim: int = 1024
    mask_res: int = 28
    width: float = 1.0                    # channel multiplier for test-size models


def make_anchors(feat_hw: list[tuple[int, int]], strides, sizes, ratios, device) -> list[torch.Tensor]:
    out = []
    for (h, w), s, size in zip(feat_hw, strides, sizes):
        r = torch.tensor(ratios, device=device, dtype=torch.float32)
        hs = size * torch.sqrt(r)
        ws = size / torch.sqrt(r)
        base = torch.stack([-ws, -hs, ws, hs], 1) / 2
        sx = (torch.arange(w, device=device, dtype=torch.float32) + 0.5) * s
        sy = (torch.arange(h, device=device, dtype=torch.float32) + 0.5) * s
        cy, cx = torch.meshgrid(sy, sx, indexing="ij")
        shifts = torch.stack([cx, cy, cx, cy], -1).reshape(-1, 1, 4)
        out.append((shifts + base[None]).reshape(-1, 4))
    return out


def encode_boxes(gt, anchors, weights):
    wx, wy, ww, wh = weights
    aw, ah = anchors[:, 2] - anchors[:, 0], anchors[:, 3] - anchors[:, 1]
    ax, ay = anchors[:, 0] + 0.5 * aw, anchors[:, 1] + 0.5 * ah
    gw, gh = (gt[:, 2] - gt[:, 0]).clamp(min=1), (gt[:, 3] - gt[:, 1]).clamp(min=1)
    gx, gy = gt[:, 0] + 0.5 * gw, gt[:, 1] + 0.5 * gh
    return torch.stack([wx * (gx - ax) / aw, wy * (gy - ay) / ah, ww * torch.log(gw / aw),
                        wh * torch.log(gh / ah)], 1)


def _sample(pos_mask, neg_mask, batch, pos_frac):
    pos, neg = pos_mask.nonzero().flatten(), neg_mask.nonzero().flatten()
    npos = min(int(batch * pos_frac), pos.numel())
    nneg = min(batch - npos, neg.numel())
    pos = pos[torch.randperm(pos.numel(), device=pos.device)[:npos]]
    neg = neg[torch.randperm(neg.numel(), device=neg.device)[:nneg]]
    return pos, neg


def _sample_static(pos_mask, neg_mask, batch: int, pos_frac: float):
    """Fixed-shape version of the fg / bg sampler: no `nonzero`, no data-dependent sizes, so the step never waits for
    the GPU.  Masks are [n] or [N, n] (one row per image, sampled independently).  Returns (pos_idx [.., P],
    pos_ok [.., P], neg_idx [.., Q], neg_ok [.., Q]) with P = batch * pos_frac, Q = batch: random members of each set
    (top-k of uniform noise restricted to the set), `*_ok` marking the entries that are real and inside the quota -
    negatives fill whatever the positives leave of `batch`, as in tensorpack's sampler."""
    P = int(batch * pos_frac)
    n = pos_mask.shape[-1]
    noise = torch.rand((2, *pos_mask.shape), device=pos_mask.device)
    ps, pos_idx = torch.where(pos_mask, noise[0], noise.new_full((), -1.0)).topk(min(P, n), dim=-1)
    pos_ok = ps >= 0
    quota = batch - pos_ok.sum(-1, keepdim=True)                   # device values: never read on the host
    ns, neg_idx = torch.where(neg_mask, noise[1], noise.new_full((), -1.0)).topk(min(batch, n), dim=-1)
    neg_ok = (ns >= 0) & (torch.arange(neg_idx.shape[-1], device=ns.device) < quota)
    return pos_idx, pos_ok, neg_idx, neg_ok


def pad_targets(targets):
    """Per-image target lists -> ([N, G, 4] boxes, [N, G] labels, [N, G] valid) with G = the largest count of the
    batch (at least 1); padding boxes are zeros, which have IoU 0 with everything.  Shapes come from the host-side
    sizes of the lists, values never leave the device."""
    from torch.nn.utils.rnn import pad_sequence

    dev = targets[0]["boxes"].device
    boxes = pad_sequence([t["boxes"].float().reshape(-1, 4) for t in targets], batch_first=True)
    labels = pad_sequence([t["labels"].long().reshape(-1) for t in targets], batch_first=True)
    valid = pad_sequence([torch.ones(t["boxes"].shape[0], dtype=torch.bool, device=dev) for t in targets], batch_first=True)
    if boxes.shape[1] == 0:
        n = len(targets)
        boxes, labels = boxes.new_zeros(n, 1, 4), labels.new_zeros(n, 1)
        valid = torch.zeros(n, 1, dtype=torch.bool, device=dev)
    return boxes, labels, valid


def _masked(v, ok):
    """v where `ok` (per row), exact zeros elsewhere - also when v is inf or nan there."""
    ok = ok.view(-1, *([1] * (v.dim() - 1)))
    return torch.where(ok, v, torch.zeros_like(v))


class RPNHead(nn.Module):
    def __init__(self, c, a):
        super().__init__()
        self.conv = nn.Conv2d(c, c, 3, padding=1)
        self.cls = nn.Conv2d(c, a, 1)
        self.box = nn.Conv2d(c, a * 4, 1)
        for m in (self.conv, self.cls, self.box):
            nn.init.normal_(m.weight, std=0.01)
            nn.init.zeros_(m.bias)

    def forward(self, feats):
        logits, deltas = [], []
        for f in feats:
            t = D.conv_bias_act(f, self.conv, relu=True)
            logits.append(self.cls(t).permute(0, 2, 3, 1).flatten(1))               # [N, H*W*A]
            n, _, h, w = f.shape
            deltas.append(self.box(t).view(n, -1, 4, h, w).permute(0, 3, 4, 1, 2).reshape(n, -1, 4))
        return logits, deltas


class FCHead(nn.Module):
    """2-fc box head + class / box predictors on the tcgen05 GEMM."""

    def __init__(self, in_dim, fc, ncls):
        super().__init__()
        mk = lambda o, i, std: nn.Parameter(torch.empty(o, i).normal_(0, std))  # noqa: E731
        self.w6, self.b6 = mk(fc, in_dim, math.sqrt(2.0 / in_dim)), nn.Parameter(torch.zeros(fc))
        self.w7, self.b7 = mk(fc, fc, math.sqrt(2.0 / fc)), nn.Parameter(torch.zeros(fc))
        # predictor widths are padded to a multiple of 8 (16-byte rows for the TMA-fed GEMMs)
        self.ncls = ncls
        pc, pb = (ncls + 7) // 8 * 8, (ncls * 4 + 7) // 8 * 8
        self.wc, self.bc = mk(pc, fc, 0.01), nn.Parameter(torch.zeros(pc))
        self.wb, self.bb = mk(pb, fc, 0.001), nn.Parameter(torch.zeros(pb))

    def forward(self, x):
        x = F.relu(G.linear(x, self.w6, self.b6))
        x = F.relu(G.linear(x, self.w7, self.b7))
        return (G.linear(x, self.wc, self.bc)[:, :self.ncls].float(),
                G.linear(x, self.wb, self.bb)[:, :self.ncls * 4].float())


class MaskHead(nn.Module):
    def __init__(self, c, ncls, dim=256):
        super().__init__()
        self.convs = nn.ModuleList([nn.Conv2d(c if i == 0 else dim, dim, 3, padding=1) for i in range(4)])
        self.deconv = nn.ConvTranspose2d(dim, dim, 2, stride=2)
        self.pred = nn.Conv2d(dim, ncls, 1)
        for m in [*self.convs, self.deconv, self.pred]:
            nn.init.kaiming_normal_(m.weight, mode="fan_out", nonlinearity="relu")
            nn.init.zeros_(m.bias)

    def forward(self, x):
        for c in self.convs:
            x = D.conv_bias_act(x, c, relu=True)
        return self.pred(F.relu(self.deconv(x)))


class MaskRCNN(nn.Module):
    def __init__(self, cfg: MaskRCNNConfig | None = None):
        super().__init__()
        from .resnet_fpn import ResNetFPN

        self.cfg = cfg = cfg or MaskRCNNConfig()
        if not cfg.mode_fpn:
            raise NotImplementedError("MODE_FPN=False (tensorpack's ResNet-C4 variant) is not built; every Mask R-CNN "
                                      "configuration of the reference sets MODE_FPN=True")
        c = max(16, int(256 * cfg.width) // 8 * 8)
        # ResNet-50 (ResNet-18 for the test-size models) + FPN in repo code: BACKBONE.NORM=FreezeBN is folded into the
        # convolution weights (models/resnet_fpn.py), the stem and res2 are frozen as tensorpack does
        self.backbone = ResNetFPN(50 if cfg.width >= 1.0 else 18, c)
        self.strides = (4, 8, 16, 32, 64)
        # fixed-shape training path (no host synchronisation inside the step); B2F_MASKRCNN_STATIC=0 restores the
        # variable-length formulation, which is also what the CPU oracle tests compare against
        import os

        self.static_shapes = os.environ.get("B2F_MASKRCNN_STATIC", "1") == "1"
        self.rpn = RPNHead(c, len(cfg.anchor_ratios))
        self.box_head = FCHead(c * 7 * 7, max(64, int(cfg.fc_dim * cfg.width)), cfg.num_classes)
        self.mask_head = MaskHead(c, cfg.num_classes, c) if cfg.mode_mask else None

    # ------------------------------------------------------------------ trunk: backbone + FPN + RPN head
    def _run_trunk(self, x, training):
        g = self.__dict__.get("_trunk_graph")
        if g is not None and training and tuple(x.shape) == g[1] and x.dtype == g[2]:
            out = g[0](x)
            return list(out[:5]), list(out[5:10]), list(out[10:15])
        fmap = self.backbone(x)
        feats = [fmap[k] for k in ("0", "1", "2", "3", "pool")]
        logits, deltas = self.rpn(feats)
        return feats, logits, deltas

    def enable_trunk_graph(self, example_images: torch.Tensor):
        """Capture backbone + FPN + RPN head (forward and backward) as CUDA graphs for this image shape: the part of
        the training step whose shapes never depend on the data, and more than half of its kernel launches.  Frozen-BN
        folds of the frozen layers are baked in, so call this after the backbone weights are loaded; other shapes,
        eval mode and CPU keep the eager path."""
        if not example_images.is_cuda:
            return False

        class Trunk(nn.Module):
            def __init__(self, backbone, rpn):
                super().__init__()
                self.backbone, self.rpn = backbone, rpn

            def forward(self, x):
                fmap = self.backbone(x)
                feats = [fmap[k] for k in ("0", "1", "2", "3", "pool")]
                logits, deltas = self.rpn(feats)
                return (*feats, *logits, *deltas)

        trunk = Trunk(self.backbone, self.rpn)
        trunk.train(self.training)
        x = example_images.detach().contiguous(memory_format=torch.channels_last).clone()
        graphed = torch.cuda.make_graphed_callables(trunk, (x,))
        self.__dict__["_trunk_graph"] = (graphed, tuple(x.shape), x.dtype)     # not a registered submodule
        return True

    # ------------------------------------------------------------------ RPN
    def _proposals(self, logits, deltas, anchors, img_hw, training, static=False):
        cfg = self.cfg
        n = logits[0].shape[0]
        pre = cfg.rpn_pre_nms_train if training else cfg.rpn_pre_nms_test
        post = cfg.rpn_post_nms_train if training else cfg.rpn_post_nms_test
        boxes_all, scores_all, seg_sizes = [], [], []
        for lg, dl, an in zip(logits, deltas, anchors):
            k = min(pre, lg.shape[1])
            sc, idx = lg.detach().float().topk(k, dim=1)                     # sorted descending
            d = torch.gather(dl.detach().float(), 1, idx[..., None].expand(-1, -1, 4))
            a = an[idx]
            b = D.box_decode(a.reshape(-1, 4), d.reshape(-1, 4), (1.0, 1.0, 1.0, 1.0), img_hw[1], img_hw[0])
            boxes_all.append(b.view(n, k, 4))
            scores_all.append(sc)
            seg_sizes.append(k)
        # segments ordered (image, level): each is already sorted by score
        boxes = torch.cat([torch.cat([b[i] for b in boxes_all], 0) for i in range(n)], 0)
        scores = torch.cat([torch.cat([s[i] for s in scores_all], 0) for i in range(n)], 0)
        sizes = torch.tensor(seg_sizes * n, device=boxes.device)
        seg_off = torch.cat([sizes.new_zeros(1), sizes.cumsum(0)]).int()
        keep = D.nms_batched(boxes, seg_off, cfg.rpn_nms_thresh, max_seg=max(seg_sizes))
        wh_ok = ((boxes[:, 2] - boxes[:, 0]) > 1e-3) & ((boxes[:, 3] - boxes[:, 1]) > 1e-3)
        scores = torch.where(keep & wh_ok, scores, scores.new_full((), -1e9))
        per_img = sum(seg_sizes)
        k = min(post, per_img)
        top, idx = scores.view(n, per_img).topk(k, dim=1)
        picked = torch.gather(boxes.view(n, per_img, 4), 1, idx[..., None].expand(-1, -1, 4))
        if static:      # fixed [n, k, 4] + validity mask: nothing on the host depends on how many survived NMS
            return picked, top > -1e8
        return [picked[i][top[i] > -1e8] for i in range(n)]

    def _rpn_loss(self, logits, deltas, anchors, targets):
        cfg = self.cfg
        lg = torch.cat(logits, 1).float()
        dl = torch.cat(deltas, 1).float()
        an = torch.cat(anchors, 0)
        cls_loss, box_loss = lg.new_zeros(()), lg.new_zeros(())
        for i, t in enumerate(targets):
            gt = t["boxes"].float()
            if gt.numel() == 0:
                labels = torch.zeros(an.shape[0], device=an.device)
                pos, neg = _sample(labels > 1, labels == 0, cfg.rpn_batch, 0.5)
            else:
                iou = D.pairwise_iou(an, gt)
                best, arg = iou.max(1)
                labels = torch.full_like(best, -1)
                labels[best < cfg.rpn_bg_iou] = 0
                labels[best >= cfg.rpn_fg_iou] = 1
                labels[iou.argmax(0)] = 1                                   # every gt gets an anchor
                pos, neg = _sample(labels == 1, labels == 0, cfg.rpn_batch, 0.5)
                if pos.numel():
                    tgt = encode_boxes(gt[arg[pos]], an[pos], (1.0, 1.0, 1.0, 1.0))
                    box_loss = box_loss + F.smooth_l1_loss(dl[i, pos], tgt, beta=1.0 / 9, reduction="sum") / cfg.rpn_batch
            sel = torch.cat([pos, neg])
            y = torch.cat([torch.ones_like(pos, dtype=torch.float32), torch.zeros_like(neg, dtype=torch.float32)])
            cls_loss = cls_loss + F.binary_cross_entropy_with_logits(lg[i, sel], y, reduction="sum") / cfg.rpn_batch
        n = max(1, len(targets))
        return cls_loss / n, box_loss / n

    def _rpn_loss_static(self, logits, deltas, anchors, gt_boxes, gt_valid):
        """`_rpn_loss` with fixed shapes and all images at once (see `_sample_static`, `pad_targets`): the same labels
        and the same losses in expectation, one pass of launches for the whole batch, no host synchronisation."""
        cfg = self.cfg
        lg = torch.cat(logits, 1).float()                               # [N, A]
        dl = torch.cat(deltas, 1).float()                               # [N, A, 4]
        an = torch.cat(anchors, 0)                                      # [A, 4]
        N, G = gt_valid.shape
        A = an.shape[0]
        iou = D.pairwise_iou(an, gt_boxes.reshape(N * G, 4)).view(A, N, G).permute(1, 0, 2)     # [N, A, G]
        iou = torch.where(gt_valid[:, None, :], iou, iou.new_full((), -1.0))
        best, arg = iou.max(2)                                          # [N, A]
        labels = torch.full_like(best, -1)
        labels = torch.where(best < cfg.rpn_bg_iou, torch.zeros_like(labels), labels)
        labels = torch.where(best >= cfg.rpn_fg_iou, torch.ones_like(labels), labels)
        top_anchor = iou.argmax(1)                                      # [N, G]: every valid gt gets an anchor
        labels.scatter_(1, top_anchor, torch.where(gt_valid, torch.ones_like(best[:, :G]),
                                                    torch.gather(labels, 1, top_anchor)))
        pi, pok, ni, nok = _sample_static(labels == 1, labels == 0, cfg.rpn_batch, 0.5)        # [N, P], [N, Q]
        gsel = torch.gather(arg, 1, pi)                                 # [N, P] gt of each sampled positive
        gtb = torch.gather(gt_boxes, 1, gsel[..., None].expand(-1, -1, 4))
        # unused slots are selected away, not multiplied away: 0 * inf would poison the sum
        tgt = _masked(encode_boxes(gtb.reshape(-1, 4), an[pi.reshape(-1)], (1.0, 1.0, 1.0, 1.0)), pok.reshape(-1))
        pred = torch.gather(dl, 1, pi[..., None].expand(-1, -1, 4)).reshape(-1, 4)
        bl = F.smooth_l1_loss(pred, tgt, beta=1.0 / 9, reduction="none").sum(1)
        box_loss = _masked(bl, pok.reshape(-1)).sum() / cfg.rpn_batch
        sel = torch.cat([pi, ni], 1)
        y = torch.cat([torch.ones_like(pi, dtype=torch.float32), torch.zeros_like(ni, dtype=torch.float32)], 1)
        ce = F.binary_cross_entropy_with_logits(torch.gather(lg, 1, sel), y, reduction="none")
        cls_loss = _masked(ce.reshape(-1), torch.cat([pok, nok], 1).reshape(-1)).sum() / cfg.rpn_batch
        return cls_loss / max(1, N), box_loss / max(1, N)

    def _train_static(self, feats_nhwc, dt, logits, deltas, anchors, targets, props, pvalid, images):
        """The training half of `forward` with fixed shapes, all images at once: per image `roi_batch` sampled RoIs (a
        quarter of them foreground slots, unused slots masked out of the losses) and `roi_batch * roi_fg_fraction`
        mask RoIs."""
        cfg = self.cfg
        dev = images.device
        gt_boxes, gt_labels, gt_valid = pad_targets(targets)
        N, G = gt_valid.shape
        rpn_cls, rpn_box = self._rpn_loss_static(logits, deltas, anchors, gt_boxes, gt_valid)
        p = torch.cat([props, gt_boxes], 1)                                       # gt boxes join the proposals
        valid = torch.cat([pvalid, gt_valid], 1)                                  # [N, K]
        K = p.shape[1]
        pb = torch.arange(N, device=dev, dtype=torch.int32)[:, None].expand(N, K)
        gt_off = torch.arange(N + 1, device=dev, dtype=torch.int32) * G
        # padding boxes have zero area: IoU 0 with every proposal, so they are never anybody's foreground match
        iou, idx = D.match_proposals(p.reshape(-1, 4), pb.reshape(-1), gt_boxes.reshape(-1, 4), gt_off)
        iou, idx = iou.view(N, K), idx.view(N, K)
        fi, fok, bi, bok = _sample_static(valid & (iou >= cfg.roi_fg_iou), valid & (iou < cfg.roi_fg_iou),
                                          cfg.roi_batch, cfg.roi_fg_fraction)     # [N, Fg], [N, Bg]
        Fg = fi.shape[1]
        sel = torch.cat([fi, bi], 1)
        ok = torch.cat([fok, bok], 1).reshape(-1)
        gidx = torch.gather(idx, 1, fi).long().clamp(0, N * G - 1)                # [N, Fg] global gt index
        fg_labels = torch.where(fok, gt_labels.reshape(-1)[gidx], torch.zeros_like(gidx))
        labels = torch.cat([fg_labels, torch.zeros_like(bi)], 1).reshape(-1)
        boxes_sel = torch.gather(p, 1, sel[..., None].expand(-1, -1, 4))          # [N, Fg + Bg, 4]
        bcol = torch.arange(N, device=dev, dtype=torch.float32)[:, None, None].expand(N, sel.shape[1], 1)
        rois = torch.cat([bcol, boxes_sel], 2).reshape(-1, 5)
        fg_boxes = boxes_sel[:, :Fg]
        reg_fg = _masked(encode_boxes(gt_boxes.reshape(-1, 4)[gidx.reshape(-1)], fg_boxes.reshape(-1, 4),
                                      cfg.box_weights), fok.reshape(-1)).view(N, Fg, 4)
        reg_t = torch.cat([reg_fg, reg_fg.new_zeros(N, bi.shape[1], 4)], 1).reshape(-1, 4)
        fgw = torch.cat([fok, torch.zeros_like(bok)], 1).reshape(-1)
        box_feat = self._roi_feats(feats_nhwc, rois, 7).permute(0, 3, 1, 2).flatten(1).to(dt)
        cls_logits, box_reg = self.box_head(box_feat)
        denom = ok.sum().clamp(min=1).float()
        loss_cls = _masked(F.cross_entropy(cls_logits, labels, reduction="none"), ok).sum() / denom
        br = box_reg.view(-1, cfg.num_classes, 4)[torch.arange(labels.numel(), device=dev), labels]
        loss_box = _masked(F.smooth_l1_loss(br, reg_t, beta=1.0, reduction="none").sum(1), fgw).sum() / denom
        losses = {"rpn_cls": rpn_cls, "rpn_box": rpn_box, "frcnn_cls": loss_cls, "frcnn_box": loss_box}
        counts = [t["boxes"].shape[0] for t in targets]                            # host-side sizes
        if cfg.mode_mask and sum(counts):
            # masks stay unpadded: [sum of counts, H, W]; the padded gt index (image * G + local) becomes an index
            # into that stack through the per-image offsets
            offs = [0]
            for c in counts:
                offs.append(offs[-1] + c)
            offs_dev = torch.tensor(offs[:-1], device=dev)
            midx = (offs_dev[gidx // G] + gidx % G).clamp(max=offs[-1] - 1)
            all_masks = torch.cat([t["masks"] for t in targets if t["masks"].shape[0]], 0)
            m_rois = torch.cat([bcol[:, :Fg], fg_boxes], 2).reshape(-1, 5)
            m_tgt = D.mask_crop(all_masks, fg_boxes.reshape(-1, 4), midx.reshape(-1).int(), cfg.mask_res)
            mfeat = self._roi_feats(feats_nhwc, m_rois, 14).permute(0, 3, 1, 2).to(dt)
            mlogits = self.mask_head(mfeat.contiguous(memory_format=torch.channels_last)).float()
            ml = fg_labels.reshape(-1)
            sel_logits = mlogits[torch.arange(ml.numel(), device=dev), ml]
            per_roi = F.binary_cross_entropy_with_logits(sel_logits, m_tgt, reduction="none").mean((1, 2))
            mw = fok.reshape(-1)
            losses["mrcnn"] = _masked(per_roi, mw).sum() / mw.sum().clamp(min=1).float()
        losses["total"] = sum(losses.values())
        return losses

    # ------------------------------------------------------------- ROI heads
    def _roi_feats(self, feats_nhwc, rois, P):
        scales = [1.0 / s for s in self.strides[:4]]
        return D.roi_align(feats_nhwc[:4], scales, rois, P, sampling=2, aligned=True)   # [R, P, P, C]

    def forward(self, images: torch.Tensor, targets: list[dict] | None = None):
        """images: [N, 3, H, W]; targets: per image {"boxes" [n,4], "labels" [n], "masks" u8 [n,H,W]}.
        Training returns a loss dict; eval returns per-image detections."""
        cfg = self.cfg
        training = self.training and targets is not None
        x = images.contiguous(memory_format=torch.channels_last)
        feats, logits, deltas = self._run_trunk(x, training)
        img_hw = images.shape[-2:]
        anchors = make_anchors([f.shape[-2:] for f in feats], self.strides, cfg.anchor_sizes,
                               cfg.anchor_ratios, images.device)
        static = training and self.static_shapes
        proposals = self._proposals(logits, deltas, anchors, img_hw, training, static=static)
        feats_nhwc = [f.permute(0, 2, 3, 1).contiguous() if not f.permute(0, 2, 3, 1).is_contiguous()
                      else f.permute(0, 2, 3, 1) for f in feats]
        feats_nhwc = [f.to(torch.bfloat16) if f.is_cuda else f for f in feats_nhwc]
        dt = feats[0].dtype
        if static:
            return self._train_static(feats_nhwc, dt, logits, deltas, anchors, targets, proposals[0], proposals[1], images)
        if training:
            rpn_cls, rpn_box = self._rpn_loss(logits, deltas, anchors, targets)
            rois, labels, reg_t, mask_rois, mask_gt, mask_lbl = [], [], [], [], [], []
            gt_all = torch.cat([t["boxes"].float() for t in targets], 0)
            gt_off = torch.tensor([0] + [t["boxes"].shape[0] for t in targets], device=images.device).cumsum(0).int()
            for i, (p, t) in enumerate(zip(proposals, targets)):
                p = torch.cat([p, t["boxes"].float()], 0)                     # gt boxes join the proposals
                pb = torch.full((p.shape[0],), i, device=p.device, dtype=torch.int32)
                iou, idx = D.match_proposals(p, pb, gt_all, gt_off)
                fg, bg = _sample(iou >= cfg.roi_fg_iou, iou < cfg.roi_fg_iou, cfg.roi_batch, cfg.roi_fg_fraction)
                sel = torch.cat([fg, bg])
                lab = torch.zeros(sel.numel(), dtype=torch.long, device=p.device)
                gidx = idx[fg].long()
                all_labels = torch.cat([tt["labels"] for tt in targets])
                lab[:fg.numel()] = all_labels[gidx]
                rois.append(torch.cat([pb[sel].float()[:, None], p[sel]], 1))
                labels.append(lab)
                reg = torch.zeros(sel.numel(), 4, device=p.device)
                if fg.numel():
                    reg[:fg.numel()] = encode_boxes(gt_all[gidx], p[fg], cfg.box_weights)
                reg_t.append(reg)
                if cfg.mode_mask and fg.numel():
                    mask_rois.append(torch.cat([pb[fg].float()[:, None], p[fg]], 1))
                    mask_gt.append((i, gidx - gt_off[i].long()))
                    mask_lbl.append(all_labels[gidx])
            rois, labels, reg_t = torch.cat(rois), torch.cat(labels), torch.cat(reg_t)
            box_feat = self._roi_feats(feats_nhwc, rois, 7).permute(0, 3, 1, 2).flatten(1).to(dt)
            cls_logits, box_reg = self.box_head(box_feat)
            loss_cls = F.cross_entropy(cls_logits, labels)
            fg = (labels > 0).nonzero().flatten()
            br = box_reg.view(-1, cfg.num_classes, 4)[fg, labels[fg]]
            loss_box = F.smooth_l1_loss(br, reg_t[fg], beta=1.0, reduction="sum") / max(1, labels.numel())
            losses = {"rpn_cls": rpn_cls, "rpn_box": rpn_box, "frcnn_cls": loss_cls, "frcnn_box": loss_box}
            if cfg.mode_mask and mask_rois:
                mr = torch.cat(mask_rois)
                mfeat = self._roi_feats(feats_nhwc, mr, 14).permute(0, 3, 1, 2).to(dt)
                mlogits = self.mask_head(mfeat.contiguous(memory_format=torch.channels_last)).float()
                tgt, off = [], 0
                for (i, gi), r in zip(mask_gt, mask_rois):
                    tgt.append(D.mask_crop(targets[i]["masks"], r[:, 1:], gi.int(), cfg.mask_res))
                ml = torch.cat(mask_lbl)
                sel_logits = mlogits[torch.arange(ml.numel(), device=ml.device), ml]
                losses["mrcnn"] = F.binary_cross_entropy_with_logits(sel_logits, torch.cat(tgt))
            losses["total"] = sum(losses.values())
            return losses
        # ---------------------------------------------------------- inference
        results = []
        for i, p in enumerate(proposals):
            if p.numel() == 0:
                results.append({"boxes": p.new_zeros(0, 4), "scores": p.new_zeros(0), "labels": p.new_zeros(0).long()})
                continue
            rois = torch.cat([p.new_full((p.shape[0], 1), i), p], 1)
            feat = self._roi_feats(feats_nhwc, rois, 7).permute(0, 3, 1, 2).flatten(1).to(dt)
            cls_logits, box_reg = self.box_head(feat)
            prob = cls_logits.softmax(-1)[:, 1:]                              # drop background
            R, K = prob.shape
            boxes = D.box_decode(p[:, None].expand(-1, K, -1).reshape(-1, 4),
                                 box_reg.view(R, cfg.num_classes, 4)[:, 1:].reshape(-1, 4),
                                 cfg.box_weights, img_hw[1], img_hw[0]).view(R, K, 4)
            sc, cl = prob.flatten(), torch.arange(1, K + 1, device=p.device).repeat(R)
            bx = boxes.reshape(-1, 4)
            ok = (sc > cfg.det_score_thresh).nonzero().flatten()
            sc, cl, bx = sc[ok], cl[ok], bx[ok]
            # per-class NMS: sort by (class, -score) and cut one segment per class
            order = torch.argsort(cl.float() * 2 - sc)       # class ascending, score descending inside
            sc, cl, bx = sc[order], cl[order], bx[order]
            counts = torch.bincount(cl, minlength=K + 1)
            seg_off = torch.cat([counts.new_zeros(1), counts.cumsum(0)]).int()
            keep = D.nms_batched(bx, seg_off, cfg.det_nms_thresh) if bx.numel() else torch.zeros(0, dtype=torch.bool, device=p.device)
            sc, cl, bx = sc[keep], cl[keep], bx[keep]
            top = sc.argsort(descending=True)[:cfg.det_per_image]
            det = {"boxes": bx[top], "scores": sc[top], "labels": cl[top], "fastrcnn_all_scores": prob}
            if cfg.mode_mask and top.numel():
                mr = torch.cat([bx.new_full((top.numel(), 1), i), bx[top]], 1)
                mfeat = self._roi_feats(feats_nhwc, mr, 14).permute(0, 3, 1, 2).to(dt)
                ml = self.mask_head(mfeat.contiguous(memory_format=torch.channels_last)).float().sigmoid()
                det["masks"] = ml[torch.arange(top.numel(), device=ml.device), cl[top]]   # [n, 28, 28]
            results.append(det)
        return results


def tensorpack_to_torchvision_names(blocks=(3, 4, 6, 3)) -> dict[str, tuple[str, str]]:
    """tensorpack ResNet variable name -> (torchvision parameter / buffer name, kind) for the body of a
    bottleneck ResNet.  Kinds: "conv" (HWIO -> OIHW), "vec".  The reference starts Mask R-CNN from
    `BACKBONE.WEIGHTS=.../ImageNet-R50-AlignPadding.npz` (examples/maskrcnn/train-maskrcnn-tensorpack.yaml:32,
    fetched by charts/machine-learning/data-prep/coco-data/templates/coco-data.yaml:31-35)."""
    bn = {"gamma": "weight", "beta": "bias", "mean/EMA": "running_mean", "variance/EMA": "running_var"}
    out = {"conv0/W": ("conv1.weight", "conv")}
    for k, v in bn.items():
        out[f"conv0/bn/{k}"] = (f"bn1.{v}", "vec")
    for g, nb in enumerate(blocks):
        for b in range(nb):
            tp, tv = f"group{g}/block{b}", f"layer{g + 1}.{b}"
            for i in (1, 2, 3):
                out[f"{tp}/conv{i}/W"] = (f"{tv}.conv{i}.weight", "conv")
                for k, v in bn.items():
                    out[f"{tp}/conv{i}/bn/{k}"] = (f"{tv}.bn{i}.{v}", "vec")
            if b == 0:
                out[f"{tp}/convshortcut/W"] = (f"{tv}.downsample.0.weight", "conv")
                for k, v in bn.items():
                    out[f"{tp}/convshortcut/bn/{k}"] = (f"{tv}.downsample.1.{v}", "vec")
    return out


def load_tensorpack_backbone(body: nn.Module, npz_path) -> dict:
    """Copy a tensorpack ResNet `.npz` into a torchvision-style body.  Returns a report
    {"loaded": n, "missing": [...], "unused": [...], "placeholder": bool}; the offline stand-in written by
    the coco-data chart (`synthetic` key only) is reported as a placeholder and leaves the body as it is."""
    import numpy as np

    with np.load(str(npz_path)) as z:
        keys = set(z.files)
        if keys <= {"synthetic"}:
            return {"loaded": 0, "missing": [], "unused": [], "placeholder": True}
        table = tensorpack_to_torchvision_names()
        sd = body.state_dict()
        loaded, missing = 0, []
        for tp_name, (tv_name, kind) in table.items():
            if tp_name not in keys or tv_name not in sd:
                missing.append(tp_name)
                continue
            w = torch.from_numpy(np.asarray(z[tp_name]))
            if kind == "conv":
                w = w.permute(3, 2, 0, 1)          # [kh, kw, in, out] -> [out, in, kh, kw]
            assert tuple(w.shape) == tuple(sd[tv_name].shape), (tp_name, tuple(w.shape), tuple(sd[tv_name].shape))
            sd[tv_name].copy_(w.to(sd[tv_name].dtype))
            loaded += 1
        unused = sorted(k for k in keys if k not in table)
    return {"loaded": loaded, "missing": missing, "unused": unused, "placeholder": False}


def paste_masks(det: dict, H: int, W: int) -> torch.Tensor:
    """28x28 probabilities -> full-image binary masks (the notebook's `get_mask` step)."""
    if "masks" not in det or det["masks"].numel() == 0:
        return torch.zeros(0, H, W, dtype=torch.uint8)
    return D.mask_paste(det["masks"], det["boxes"], H, W, 0.5)
