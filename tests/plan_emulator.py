"""TEST INFRASTRUCTURE: a torch-CPU stand-in for engine.Context that records the planned ops and
executes them with torch in fp32 on NHWC channel slices.  It lets the CPU suite check the host-side
graph wiring (buffer slices, merged cv1||cv2, in-place residual, space-to-depth stem, weight packing)
against the oracle without a GPU.  It is never imported by the product."""
from __future__ import annotations

import numpy as np
import torch
import torch.nn.functional as F


class PlanEmulator:
    def __init__(self):
        self.bufs = []
        self.ops = []
        self.batch = None

    # --- the subset of engine.Context that TileDetector._build uses
    def plan_begin(self, batch, net_h, net_w, nc):
        self.batch, self.net_h, self.net_w, self.nc = batch, net_h, net_w, nc
        self.rows = 3 * sum((net_h // s) * (net_w // s) for s in (8, 16, 32))

    def buffer(self, h, w, c):
        self.bufs.append(torch.zeros(self.batch, h, w, c))
        return len(self.bufs) - 1

    def stage_begin(self, sub_batch, lanes=1):  # depth-first stages change the launch order, not the values
        self.stage_calls = getattr(self, "stage_calls", 0) + 1

    def stage_end(self):
        pass

    def conv(self, src, dst, w_oihw, bias, k, stride, act=True, res=None, tile_n=0, mblk=0, mode=0, cluster=0, up=None):
        # weights go through the same fp16 rounding the device sees
        self.ops.append(("conv", src, dst, w_oihw.half().float(), bias.float(), k, stride, act, res, up))

    def conv_chain(self, src, dst, w_oihw, bias, k, stride, act, dst2, w2_oihw, bias2, act2, store_first=True, up=None, mblk=0, mode=0):
        # two convs in one kernel on the device: the same two ops here
        self.conv(src, dst, w_oihw, bias, k, stride, act, up=up)
        self.conv(dst, dst2, w2_oihw, bias2, 1, 1, act2)

    def detect_level(self, src, w_oi, bias, nc, level_row_off, stride, anchors_px):
        self.ops.append(("detect", src, w_oi.reshape(w_oi.shape[0], -1).half().float(), bias.float(), nc, level_row_off, stride,
                         np.asarray(anchors_px, np.float32)))

    def call(self, name, *args):
        if name == "aq_plan_sppf_pool":
            self.ops.append(("pool",) + args)
        elif name == "aq_plan_upsample2x":
            self.ops.append(("up",) + args)
        elif name == "aq_plan_num_launches":
            return len(self.ops)
        else:
            raise NotImplementedError(name)
        return 0

    def plan_end(self):
        pass

    def close(self):
        pass

    # --- execution
    @torch.no_grad()
    def run(self, x_s2d: torch.Tensor, fp16_activations=False) -> torch.Tensor:
        """x_s2d: (B, H/2, W/2, 16) -> decoded predictions (B, rows, 5+nc)."""
        q = (lambda t: t.half().float()) if fp16_activations else (lambda t: t)
        self.bufs[0].copy_(x_s2d)
        pred = torch.zeros(self.batch, self.rows, self.nc + 5)
        for op in self.ops:
            if op[0] == "conv":
                _, src, dst, w, b, k, s, act, res, up = op
                c_in, c_out = w.shape[1], w.shape[0]
                if up is not None:  # input = [upsample2x(up buffer slice) | src slice]
                    u = self.bufs[up[0]][..., up[1]:up[1] + up[2]].repeat_interleave(2, 1).repeat_interleave(2, 2)
                    x = torch.cat((u, self.bufs[src[0]][..., src[1]:src[1] + c_in - up[2]]), -1).permute(0, 3, 1, 2)
                else:
                    x = self.bufs[src[0]][..., src[1]:src[1] + c_in].permute(0, 3, 1, 2)
                y = F.conv2d(x, w, b, s, k // 2)
                if act:
                    y = F.silu(y)
                y = y.permute(0, 2, 3, 1)
                if res is not None:
                    y = y + self.bufs[res[0]][..., res[1]:res[1] + c_out]
                self.bufs[dst[0]][..., dst[1]:dst[1] + c_out] = q(y)
            elif op[0] == "pool":
                _, buf, c_off, c = op
                t = self.bufs[buf]
                x = t[..., c_off:c_off + c].permute(0, 3, 1, 2)
                for i in range(1, 4):
                    x = F.max_pool2d(x, 5, 1, 2)
                    t[..., c_off + i * c:c_off + (i + 1) * c] = x.permute(0, 2, 3, 1)
            elif op[0] == "up":
                _, src, src_off, c, dst, dst_off = op
                x = self.bufs[src][..., src_off:src_off + c]
                self.bufs[dst][..., dst_off:dst_off + c] = x.repeat_interleave(2, 1).repeat_interleave(2, 2)
            else:
                _, src, w, b, nc, row_off, stride, anc = op
                no = nc + 5
                x = self.bufs[src[0]][..., src[1]:src[1] + w.shape[1]]
                B, ny, nx, _ = x.shape
                z = (x.reshape(-1, w.shape[1]) @ w.t() + b).view(B, ny, nx, 3, no).sigmoid()
                gy, gx = torch.meshgrid(torch.arange(ny).float(), torch.arange(nx).float(), indexing="ij")
                out = z.clone()
                out[..., 0] = (z[..., 0] * 2 + (gx[None, :, :, None] - 0.5)) * stride
                out[..., 1] = (z[..., 1] * 2 + (gy[None, :, :, None] - 0.5)) * stride
                a = torch.from_numpy(anc).view(3, 2)
                out[..., 2] = (z[..., 2] * 2) ** 2 * a[:, 0]
                out[..., 3] = (z[..., 3] * 2) ** 2 * a[:, 1]
                pred[:, row_off:row_off + 3 * ny * nx] = out.permute(0, 3, 1, 2, 4).reshape(B, -1, no)
        return pred


def space_to_depth_input(x_nchw_rgb: torch.Tensor) -> torch.Tensor:
    """(B,3,H,W) float in [0,1] -> (B,H/2,W/2,16) with channel (py*2+px)*3 + c, as the preprocess kernel writes."""
    B, _, H, W = x_nchw_rgb.shape
    out = torch.zeros(B, H // 2, W // 2, 16)
    for py in range(2):
        for px in range(2):
            for c in range(3):
                out[..., (py * 2 + px) * 3 + c] = x_nchw_rgb[:, c, py::2, px::2]
    return out
