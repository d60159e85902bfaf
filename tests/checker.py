"""Picks the strongest CPU checker available and gives it one interface.

Preference: the compiled reference (oracle/_ref/libdaala_ref.so, built where
/root/reference is mounted and shipped prebuilt to the GPU box), else the C
restatement (oracle/libtdlt_oracle.so).  Test infrastructure only.
"""
import numpy as np

import oracle_libs as ol


class Checker:
    def __init__(self, pic_w, pic_h, prefer_ref=True):
        self.pic_w, self.pic_h = pic_w, pic_h
        self.fw, self.fh = ol.padded_dims(pic_w, pic_h)
        self.nhsb, self.nvsb = self.fw // 32, self.fh // 32
        self.rc = None
        if prefer_ref and ol.have_ref():
            self.rc = ol.RefCtx(pic_w, pic_h)
            assert (self.rc.fw, self.rc.fh) == (self.fw, self.fh)
        self.kind = "reference" if self.rc else "port"

    def close(self):
        if self.rc:
            self.rc.close()

    def fwd_plane(self, c, bsize, dec, keyframe, haar=0):
        if self.rc:
            return self.rc.fwd_plane(c, bsize, dec, keyframe, use_haar=haar, impl=1)[0]
        return ol.port_fwd_plane(c, bsize, dec, self.pic_w, self.pic_h, keyframe, use_haar=haar)[0]

    def inv_plane(self, d, bsize, dec, dc_split, haar=0):
        if self.rc:
            return self.rc.inv_plane(d, bsize, dec, dc_split, use_haar=haar, impl=1)
        return ol.port_inv_plane(d, bsize, dec, self.pic_w, self.pic_h, dc_split, use_haar=haar)

    def cfl_pred_frames(self, d_luma, bsize):
        """d_luma: (nframes, fh, fw) -> (nframes, fh/2, fw/2) chroma-from-luma predictor planes."""
        f = ol.ref_cfl_pred_plane if self.rc else ol.port_cfl_pred_plane
        return np.stack([f(d_luma[i], bsize[i]) for i in range(d_luma.shape[0])])

    def dering_frames(self, planes, decs, flags, bskip, q):
        """planes[p]: (nframes, h, w); flags (nframes, nvsb, nhsb); bskip[p] (nframes, nvsb*8,
        nhsb*8).  The reference's od_dering superblock by superblock when it is available."""
        out = [np.empty_like(p) for p in planes]
        for f in range(flags.shape[0]):
            args = ([p[f] for p in planes], decs, flags[f], [b[f] for b in bskip], q)
            res = ol.ref_dering_frame(self.rc, *args) if self.rc else ol.port_dering_frame(*args)
            for p in range(len(planes)):
                out[p][f] = res[p]
        return out

    def smooth_frames(self, planes, decs, bsize, q):
        out = [np.empty_like(p) for p in planes]
        f = ol.ref_smooth_frame if self.rc else ol.port_smooth_frame
        for i in range(bsize.shape[0]):
            res = f([p[i] for p in planes], decs, bsize[i], q)
            for p in range(len(planes)):
                out[p][i] = res[p]
        return out

    def fwd_frames(self, c_planes, bsize, decs, keyframe, haar=0):
        """c_planes[p]: (nframes, h, w) -> list of d planes of the same shape."""
        out = []
        for p, dec in enumerate(decs):
            d = np.empty_like(c_planes[p])
            for f in range(c_planes[p].shape[0]):
                d[f] = self.fwd_plane(c_planes[p][f], bsize[f], dec, keyframe, haar)
            out.append(d)
        return out

    def inv_frames(self, d_planes, bsize, decs, dc_split, haar=0):
        out = []
        for p, dec in enumerate(decs):
            c = np.empty_like(d_planes[p])
            for f in range(d_planes[p].shape[0]):
                c[f] = self.inv_plane(d_planes[p][f], bsize[f], dec, dc_split, haar)
            out.append(c)
        return out


def make_maps(kind, nframes, nhsb, nvsb, rng):
    shape = (nframes, nvsb * 4, nhsb * 4)
    if kind == "quad":
        return np.stack([ol.random_quadtree_bsize(nhsb, nvsb, rng) for _ in range(nframes)])
    if kind == "deep":
        return np.stack([ol.random_quadtree_bsize(nhsb, nvsb, rng, 1.0, 0.9, 0.5) for _ in range(nframes)])
    if kind.startswith("uniform"):
        return np.full(shape, int(kind[-1]), np.uint8)
    raise ValueError(kind)
