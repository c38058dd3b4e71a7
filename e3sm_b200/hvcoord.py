"""Hybrid vertical coordinate (hvcoord_t, H/src/share/hybvcoord_mod.F90:19-32, 169-191).

Tables come from H/test/vcoord/*.ascii, converted once by e3sm_b200/data/make_vcoord.py into
vcoord.npz (the reference tree does not exist on the GPU box).
"""
import os
import numpy as np

from .geometry import P0

_DATA = os.path.join(os.path.dirname(os.path.abspath(__file__)), "data", "vcoord.npz")


class HVCoord:
    def __init__(self, hyai, hybi, hyam, hybm, ps0=P0):
        self.hyai = np.ascontiguousarray(hyai, dtype=np.float64)
        self.hybi = np.ascontiguousarray(hybi, dtype=np.float64)
        self.hyam = np.ascontiguousarray(hyam, dtype=np.float64)
        self.hybm = np.ascontiguousarray(hybm, dtype=np.float64)
        self.ps0 = float(ps0)
        self.nlev = self.hyam.size
        assert self.hyai.size == self.nlev + 1 and self.hybi[0] == 0.0
        # set_layer_locations hybvcoord_mod.F90:181-187
        self.etam = self.hyam + self.hybm
        self.etai = self.hyai + self.hybi
        self.dp0 = (self.hyai[1:] - self.hyai[:-1]) * self.ps0 + (self.hybi[1:] - self.hybi[:-1]) * self.ps0

    @classmethod
    def from_table(cls, nlev):
        d = np.load(_DATA)
        if f"hyai{nlev}" not in d:
            raise ValueError(f"no vertical coordinate table for nlev={nlev} (have 26, 72, 128)")
        return cls(d[f"hyai{nlev}"], d[f"hybi{nlev}"], d[f"hyam{nlev}"], d[f"hybm{nlev}"])

    def nu_scale_top(self, tom_sponge_start=0.0):
        """model_init2, H/src/theta-l/share/model_init_mod.F90:103-134 -> (nu_scale_top[nlev], nlev_tom)."""
        nst = np.zeros(self.nlev)
        nlev_tom = 0
        for k in range(self.nlev):
            if tom_sponge_start == 0:
                if self.etai[0] == 0:
                    r = self.etam[0] / self.etam[k]
                else:
                    r = self.etai[0] / self.etam[k]
                v = 16 * r ** 2 / (r ** 2 + 1)
            else:
                r = (tom_sponge_start / 1e3) / self.etam[k]
                v = 0.15 * r ** 2
            if v < 0.15:
                v = 0.0
            if v > 8.0:
                v = 8.0
            if v > 0:
                nlev_tom = k + 1
            nst[k] = v
        return nst, nlev_tom
