"""K12-K15 wrappers (csrc/kernels/detection.cu): multilevel RoIAlign, box decode, batched NMS,
proposal matcher, mask crop / paste.  CPU tensors take the torchvision / plain-torch oracle."""
from __future__ import annotations

import ctypes
import math

import torch

from . import _lib
from ._lib import c_float, c_int, c_ll, ptr, stream_ptr


def _ptr_array(tensors):
    return (ctypes.c_void_p * len(tensors))(*[t.data_ptr() for t in tensors])


def _int_array(vals):
    return (ctypes.c_int * len(vals))(*[int(v) for v in vals])


def _float_array(vals):
    return (ctypes.c_float * len(vals))(*[float(v) for v in vals])


def fpn_level_of(rois: torch.Tensor) -> torch.Tensor:
    """k = floor(4 + log2(sqrt(area) / 224)) clamped to [2, 5], returned as 0..3."""
    area = (rois[:, 3] - rois[:, 1]).clamp(min=0) * (rois[:, 4] - rois[:, 2]).clamp(min=0)
    k = torch.floor(4 + torch.log2(area.sqrt() / 224 + 1e-6))
    return (k.clamp(2, 5) - 2).long()


class _RoIAlign(torch.autograd.Function):
    @staticmethod
    def forward(ctx, rois, P, sampling, aligned, scales, *feats):
        R, C = rois.shape[0], feats[0].shape[-1]
        out = torch.empty((R, P, P, C), dtype=_lib.HALF, device=rois.device)
        H = [f.shape[1] for f in feats]
        W = [f.shape[2] for f in feats]
        feats = [f.contiguous() for f in feats]
        _lib.call("b2f_roi_align_fwd", _ptr_array(feats), _int_array(H), _int_array(W), _float_array(scales),
                  c_int(len(feats)), c_int(C), ptr(rois), c_int(R), c_int(P), c_int(sampling),
                  c_int(int(aligned)), ptr(out), stream_ptr())
        ctx.save_for_backward(rois, *feats)
        ctx.cfg = (P, sampling, aligned, scales, H, W)
        return out

    @staticmethod
    def backward(ctx, dout):
        rois, *feats = ctx.saved_tensors
        P, sampling, aligned, scales, H, W = ctx.cfg
        grads = [torch.zeros(f.shape, dtype=torch.float32, device=f.device) for f in feats]
        dout = dout.contiguous()
        _lib.call("b2f_roi_align_bwd", _ptr_array(feats), _ptr_array(grads), _int_array(H), _int_array(W),
                  _float_array(scales), c_int(len(feats)), c_int(feats[0].shape[-1]), ptr(rois),
                  c_int(rois.shape[0]), c_int(P), c_int(sampling), c_int(int(aligned)), ptr(dout), stream_ptr())
        return (None, None, None, None, None, *[g.to(_lib.HALF) for g in grads])


def roi_align(feats_nhwc: list[torch.Tensor], scales: list[float], rois: torch.Tensor, P: int,
              sampling: int = 2, aligned: bool = True) -> torch.Tensor:
    """feats: NHWC bf16 per FPN level (P2..P5); rois [R, 5]; returns [R, P, P, C]."""
    rois = rois.float().contiguous()
    if not rois.is_cuda:
        import torchvision

        C = feats_nhwc[0].shape[-1]
        out = torch.zeros((rois.shape[0], P, P, C), dtype=feats_nhwc[0].dtype)
        lv = fpn_level_of(rois) if len(feats_nhwc) > 1 else torch.zeros(rois.shape[0], dtype=torch.long)
        for i, f in enumerate(feats_nhwc):
            idx = (lv == i).nonzero().flatten()
            if idx.numel():
                o = torchvision.ops.roi_align(f.permute(0, 3, 1, 2).float(), rois[idx], P, scales[i], sampling, aligned)
                out[idx] = o.permute(0, 2, 3, 1).to(out.dtype)
        return out
    return _RoIAlign.apply(rois, P, sampling, aligned, list(scales), *feats_nhwc)


class _BiasAct(torch.autograd.Function):
    """y <- act(y + bias[c] (+ residual)) in place on a channels-last activation (`b2f_bias_act_nhwc`).  `bias` is
    either a constant (the shift of a folded frozen batch norm) or a trainable convolution bias, whose gradient is the
    column sum of the incoming gradient over N, H, W (`ops.elementwise.colsum`)."""

    @staticmethod
    def forward(ctx, y, bias, residual, relu):
        C = y.shape[1]
        _lib.call("b2f_bias_act_nhwc", ptr(y), ptr(bias), ptr(residual), c_ll(y.numel()), c_int(C), c_int(int(relu)),
                  stream_ptr())
        ctx.mark_dirty(y)
        ctx.relu, ctx.has_res, ctx.bias_dtype = relu, residual is not None, bias.dtype
        if relu:
            ctx.save_for_backward(y)
        return y

    @staticmethod
    def backward(ctx, g):
        if ctx.relu:
            (out,) = ctx.saved_tensors
            g = torch.ops.aten.threshold_backward(g, out, 0)
        db = None
        if ctx.needs_input_grad[1]:
            from .elementwise import colsum

            g = g.contiguous(memory_format=torch.channels_last)
            C = g.shape[1]
            db = colsum(g.permute(0, 2, 3, 1).reshape(-1, C)).to(ctx.bias_dtype)      # NHWC memory as [N*H*W, C]
        return g, db, (g if ctx.has_res else None), None


def bias_act_(y: torch.Tensor, bias: torch.Tensor, residual: torch.Tensor | None = None, relu: bool = True):
    """In place on the convolution output `y` [N, C, H, W]: y = act(y + bias.view(1, C, 1, 1) (+ residual))."""
    fast = (y.is_cuda and y.dtype == _lib.HALF and y.dim() == 4 and y.shape[1] % 8 == 0
            and y.is_contiguous(memory_format=torch.channels_last) and bias.dtype == y.dtype and bias.is_contiguous()
            and (residual is None or (residual.shape == y.shape and residual.dtype == y.dtype
                                      and residual.is_contiguous(memory_format=torch.channels_last))))
    if fast:
        return _BiasAct.apply(y, bias, residual, relu)
    y = y.add_(bias.view(1, -1, 1, 1).to(y.dtype))
    if residual is not None:
        y = y.add_(residual)
    return torch.relu_(y) if relu else y


def conv_bias_act(x: torch.Tensor, conv, relu: bool = False):
    """`conv` (an nn.Conv2d with a bias) followed by an optional ReLU, with the bias and the activation in one in-place
    pass over the output instead of cuDNN's separate broadcast add plus a ReLU kernel."""
    if not x.is_cuda or conv.bias is None:
        y = conv(x)
        return torch.relu_(y) if relu else y
    y = torch.nn.functional.conv2d(x, conv.weight, None, conv.stride, conv.padding, conv.dilation, conv.groups)
    return bias_act_(y, conv.bias, None, relu)

def box_decode(anchors: torch.Tensor, deltas: torch.Tensor, weights=(1.0, 1.0, 1.0, 1.0),
               img_w: float = 1e9, img_h: float = 1e9) -> torch.Tensor:
    anchors, deltas = anchors.float().contiguous(), deltas.float().contiguous()
    if not anchors.is_cuda:
        wx, wy, ww, wh = weights
        w, h = anchors[:, 2] - anchors[:, 0], anchors[:, 3] - anchors[:, 1]
        cx, cy = anchors[:, 0] + 0.5 * w, anchors[:, 1] + 0.5 * h
        dw = (deltas[:, 2] / ww).clamp(max=math.log(1000 / 16))
        dh = (deltas[:, 3] / wh).clamp(max=math.log(1000 / 16))
        pcx, pcy = deltas[:, 0] / wx * w + cx, deltas[:, 1] / wy * h + cy
        pw, ph = dw.exp() * w, dh.exp() * h
        out = torch.stack([pcx - 0.5 * pw, pcy - 0.5 * ph, pcx + 0.5 * pw, pcy + 0.5 * ph], 1)
        out[:, 0::2] = out[:, 0::2].clamp(0, img_w)
        out[:, 1::2] = out[:, 1::2].clamp(0, img_h)
        return out
    out = torch.empty_like(anchors)
    _lib.call("b2f_box_decode", ptr(anchors), ptr(deltas), ptr(out), c_ll(anchors.shape[0]),
              *[c_float(w) for w in weights], c_float(img_w), c_float(img_h), stream_ptr())
    return out


def nms_batched(boxes: torch.Tensor, seg_off: torch.Tensor, iou_thr: float, max_seg: int | None = None):
    """boxes sorted by descending score inside each segment [seg_off[s], seg_off[s+1]).
    Returns a bool keep mask.  No host synchronisation when `max_seg` is given."""
    boxes = boxes.float().contiguous()
    nseg = seg_off.numel() - 1
    if not boxes.is_cuda:
        import torchvision

        keep = torch.zeros(boxes.shape[0], dtype=torch.bool)
        so = seg_off.tolist()
        for s in range(nseg):
            b = boxes[so[s]:so[s + 1]]
            if b.numel():
                k = torchvision.ops.nms(b, torch.arange(b.shape[0], 0, -1, dtype=torch.float32), iou_thr)
                keep[so[s] + k] = True
        return keep
    if max_seg is None:
        max_seg = int((seg_off[1:] - seg_off[:-1]).max().item())
    words = (max_seg + 63) // 64
    scratch = torch.empty(max(1, boxes.shape[0] * words), dtype=torch.int64, device=boxes.device)
    keep = torch.zeros(boxes.shape[0], dtype=torch.uint8, device=boxes.device)
    _lib.call("b2f_nms_batched", ptr(boxes), ptr(seg_off.int().contiguous()), c_int(nseg), c_int(max_seg),
              c_float(iou_thr), ptr(scratch), ptr(keep), stream_ptr())
    return keep.bool()


def match_proposals(props, prop_batch, gts, gt_off):
    """Returns (best_iou [N], best_gt_index [N]) per proposal against its image's GT boxes."""
    props, gts = props.float().contiguous(), gts.float().contiguous()
    n = props.shape[0]
    if not props.is_cuda:
        import torchvision

        bi, bj = torch.zeros(n), torch.full((n,), -1, dtype=torch.int32)
        go = gt_off.tolist()
        for b in range(len(go) - 1):
            sel = (prop_batch == b).nonzero().flatten()
            if sel.numel() and go[b + 1] > go[b]:
                m = torchvision.ops.box_iou(props[sel], gts[go[b]:go[b + 1]])
                v, j = m.max(1)
                bi[sel], bj[sel] = v, (j + go[b]).int()
        return bi, bj
    bi = torch.empty(n, dtype=torch.float32, device=props.device)
    bj = torch.empty(n, dtype=torch.int32, device=props.device)
    _lib.call("b2f_match_proposals", ptr(props), ptr(prop_batch.int().contiguous()), c_ll(n), ptr(gts),
              ptr(gt_off.int().contiguous()), ptr(bi), ptr(bj), stream_ptr())
    return bi, bj


def pairwise_iou(a, b):
    a, b = a.float().contiguous(), b.float().contiguous()
    if not a.is_cuda:
        import torchvision

        return torchvision.ops.box_iou(a, b)
    out = torch.empty((a.shape[0], b.shape[0]), dtype=torch.float32, device=a.device)
    _lib.call("b2f_pairwise_iou", ptr(a), c_ll(a.shape[0]), ptr(b), c_int(b.shape[0]), ptr(out), stream_ptr())
    return out


def mask_crop_ref(masks, rois, gt_index, M):
    """Oracle of K15 crop: bilinear sample of the GT mask at the M x M bin centres, >= 0.5."""
    R = rois.shape[0]
    H, W = masks.shape[-2:]
    out = torch.zeros(R, M, M)
    for r in range(R):
        x1, y1, x2, y2 = rois[r].tolist()
        xs = x1 + (torch.arange(M) + 0.5) * (x2 - x1) / M - 0.5
        ys = y1 + (torch.arange(M) + 0.5) * (y2 - y1) / M - 0.5
        gx = (xs.clamp(0, W - 1) / max(W - 1, 1)) * 2 - 1
        gy = (ys.clamp(0, H - 1) / max(H - 1, 1)) * 2 - 1
        grid = torch.stack(torch.meshgrid(gy, gx, indexing="ij")[::-1], -1)[None]
        m = masks[int(gt_index[r])].float()[None, None]
        v = torch.nn.functional.grid_sample(m, grid, mode="bilinear", align_corners=True)[0, 0]
        valid = ((ys >= -1) & (ys <= H))[:, None] & ((xs >= -1) & (xs <= W))[None, :]
        out[r] = ((v >= 0.5) & valid).float()
    return out


def mask_crop(masks_u8, rois, gt_index, M: int = 28):
    """masks [G, H, W] uint8, rois [R, 4], gt_index [R] -> [R, M, M] float targets."""
    rois = rois.float().contiguous()
    if not rois.is_cuda:
        return mask_crop_ref(masks_u8, rois, gt_index, M)
    G, H, W = masks_u8.shape
    out = torch.empty((rois.shape[0], M, M), dtype=torch.float32, device=rois.device)
    _lib.call("b2f_mask_crop", ptr(masks_u8.contiguous()), c_int(H), c_int(W), ptr(rois),
              ptr(gt_index.int().contiguous()), c_int(rois.shape[0]), c_int(M), ptr(out), stream_ptr())
    return out


def mask_paste_ref(probs, boxes, H, W, thr=0.5):
    """Oracle of K15 paste = the notebook's arithmetic (bilinear resize of the M x M mask into the
    box, threshold) — containers/aws-samples-maskrcnn/notebooks/mask-rcnn-tensorflow-viz.ipynb:190-193."""
    R, M = probs.shape[0], probs.shape[-1]
    out = torch.zeros(R, H, W, dtype=torch.uint8)
    ys, xs = torch.arange(H) + 0.5, torch.arange(W) + 0.5
    for r in range(R):
        x1, y1, x2, y2 = boxes[r].tolist()
        bw, bh = max(x2 - x1, 1.0), max(y2 - y1, 1.0)
        mx = ((xs - x1) / bw * M - 0.5).clamp(0, M - 1)
        my = ((ys - y1) / bh * M - 0.5).clamp(0, M - 1)
        gx, gy = mx / max(M - 1, 1) * 2 - 1, my / max(M - 1, 1) * 2 - 1
        grid = torch.stack(torch.meshgrid(gy, gx, indexing="ij")[::-1], -1)[None]
        v = torch.nn.functional.grid_sample(probs[r].float()[None, None], grid, mode="bilinear",
                                            align_corners=True)[0, 0]
        inside = ((ys >= y1) & (ys <= y2))[:, None] & ((xs >= x1) & (xs <= x2))[None, :]
        out[r] = ((v > thr) & inside).to(torch.uint8)
    return out


def mask_paste(probs, boxes, H: int, W: int, thr: float = 0.5):
    probs, boxes = probs.float().contiguous(), boxes.float().contiguous()
    if not probs.is_cuda:
        return mask_paste_ref(probs, boxes, H, W, thr)
    out = torch.empty((probs.shape[0], H, W), dtype=torch.uint8, device=probs.device)
    _lib.call("b2f_mask_paste", ptr(probs), c_int(probs.shape[-1]), ptr(boxes), c_int(probs.shape[0]),
              c_int(H), c_int(W), c_float(thr), ptr(out), stream_ptr())
    return out


# --------------------------------------------------- K16: polygons and RLE (pycocotools maskApi)
def _flatten_polys(objects):
    """objects: list (per object) of lists of polygons, each [x0, y0, x1, y1, ...] (COCO `segmentation`)."""
    xy, poly_off, obj_poly = [], [0], [0]
    for polys in objects:
        for p in polys:
            pts = torch.as_tensor(p, dtype=torch.float32).reshape(-1, 2)
            xy.append(pts)
            poly_off.append(poly_off[-1] + pts.shape[0])
        obj_poly.append(len(poly_off) - 1)
    return torch.cat(xy) if xy else torch.zeros(0, 2), poly_off, obj_poly


def poly_to_mask_ref(objects, H: int, W: int) -> torch.Tensor:
    """CPU oracle: even-odd test at pixel centres, parts OR-ed.  uint8 [N, H, W]."""
    out = torch.zeros(len(objects), H, W, dtype=torch.uint8)
    px = torch.arange(W, dtype=torch.float32)[None, :] + 0.5
    py = torch.arange(H, dtype=torch.float32)[:, None] + 0.5
    for o, polys in enumerate(objects):
        inside = torch.zeros(H, W, dtype=torch.bool)
        for p in polys:
            pts = torch.as_tensor(p, dtype=torch.float32).reshape(-1, 2)
            odd = torch.zeros(H, W, dtype=torch.bool)
            n = pts.shape[0]
            for i in range(n):
                xi, yi = pts[i]
                xj, yj = pts[i - 1]
                cond = (yi > py) != (yj > py)
                if yj != yi:
                    xc = (xj - xi) * (py - yi) / (yj - yi) + xi
                    odd ^= cond & (px < xc)
            inside |= odd
        out[o] = inside.to(torch.uint8)
    return out


def poly_to_mask(objects, H: int, W: int, device=None) -> torch.Tensor:
    """COCO polygon annotations -> uint8 masks [N, H, W] (K16 on a GPU, the oracle on CPU)."""
    device = torch.device(device) if device is not None else torch.device("cuda" if torch.cuda.is_available() else "cpu")
    if device.type != "cuda":
        return poly_to_mask_ref(objects, H, W)
    xy, poly_off, obj_poly = _flatten_polys(objects)
    xy = xy.to(device).contiguous()
    po = torch.tensor(poly_off, dtype=torch.int32, device=device)
    op = torch.tensor(obj_poly, dtype=torch.int32, device=device)
    out = torch.empty(len(objects), H, W, dtype=torch.uint8, device=device)
    _lib.call("b2f_poly_to_mask", ptr(xy), ptr(po), ptr(op), c_int(len(objects)), c_int(H), c_int(W), ptr(out),
              stream_ptr())
    return out


def rle_encode(mask: torch.Tensor) -> dict:
    """COCO run-length encoding of one binary mask [H, W]: column-major runs, starting with a zero run."""
    h, w = mask.shape
    flat = mask.t().reshape(-1).to(torch.uint8).cpu()
    change = torch.nonzero(flat[1:] != flat[:-1]).squeeze(-1) + 1
    bounds = torch.cat([torch.tensor([0]), change, torch.tensor([flat.numel()])])
    counts = (bounds[1:] - bounds[:-1]).tolist()
    if flat.numel() and int(flat[0]) == 1:
        counts = [0] + counts
    return {"size": [h, w], "counts": counts}


def rle_decode(rle: dict) -> torch.Tensor:
    h, w = rle["size"]
    vals = torch.zeros(len(rle["counts"]), dtype=torch.uint8)
    vals[1::2] = 1
    flat = torch.repeat_interleave(vals, torch.tensor(rle["counts"]))
    return flat.reshape(w, h).t().contiguous()


def rle_area(rle: dict) -> int:
    return int(sum(rle["counts"][1::2]))


def rle_iou(a: dict, b: dict) -> float:
    ma, mb = rle_decode(a).bool(), rle_decode(b).bool()
    inter = int((ma & mb).sum())
    union = int((ma | mb).sum())
    return inter / union if union else 0.0
