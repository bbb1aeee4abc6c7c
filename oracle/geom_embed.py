"""CPU restatement of the reference's geometry converters (TEST INFRASTRUCTURE ONLY — never imported by the product).

Pinned to the live reference by tests/golden/ds1.npz (oracle/make_golden_ds1.py): the GeomConverter weight matrices,
NNConverter.enc / dec outputs, and DS1 denoise / DDim outputs that pass through them.

  geom_weight_mats   <- utils/utils.py:662-720  (GeomConverter.__init__)
  nn_enc / nn_dec    <- utils/utils.py:610-653  (NNConverter.enc / dec, with GeomConverter.reshape / unreshape :722-742)
  matrix_enc / _dec  <- utils/HGCal_utils.py:315-320, 340-349 (Embeder / Decoder einsums, averaging decode)
"""
import numpy as np
import torch


def geom_weight_mats(lay_r_edges):
    """Area-fraction matrices (dim_r_out x dim_in) per layer and the merged radial edges (float32, as the reference)."""
    all_edges = torch.unique(torch.tensor([e for lay in lay_r_edges for e in lay], dtype=torch.float32))
    areas = all_edges[1:] ** 2 - all_edges[:-1] ** 2
    mats = []
    for lay in lay_r_edges:
        dim_in = len(lay) - 1
        w = torch.zeros((len(all_edges) - 1, dim_in))
        for ir in range(dim_in):
            lo = int(torch.nonzero(all_edges == lay[ir])[0][0])
            hi = int(torch.nonzero(all_edges == lay[ir + 1])[0][0])
            idx = list(range(lo, hi))
            w[idx, ir] = areas[idx] / (lay[ir + 1] ** 2 - lay[ir] ** 2)
        mats.append(w)
    return mats, all_edges


def nn_enc(x, enc_w, layer_boundaries, lay_alphas, alpha_out):
    """(B, n_flat) -> (B, 1, L, alpha_out, dim_r_out): per-layer x_i W_i^T; one-alpha layers spread evenly over alpha."""
    B = x.shape[0]
    L, Ro = len(enc_w), enc_w[0].shape[0]
    out = torch.zeros((B, 1, L, alpha_out, Ro), dtype=x.dtype)
    for i in range(L):
        xi = x[:, layer_boundaries[i]:layer_boundaries[i + 1]].reshape(B, int(lay_alphas[i]), -1)
        o = xi @ enc_w[i].to(x.dtype).T
        if lay_alphas[i] == 1:
            o = torch.repeat_interleave(o, alpha_out, dim=-2) / alpha_out
        elif lay_alphas[i] != alpha_out:
            raise ValueError("Num alpha bins for layer %i is %i" % (i, lay_alphas[i]))
        out[:, 0, i] = o
    return out


def nn_dec(y, dec_w, layer_boundaries, lay_alphas, alpha_out):
    """(B, 1, L, alpha_out, dim_r_out) -> (B, n_flat): per-layer y_i D_i^T; one-alpha layers summed over alpha."""
    y = torch.squeeze(y, dim=1)
    B = y.shape[0]
    out = torch.zeros((B, int(layer_boundaries[-1])), dtype=y.dtype)
    for i in range(len(dec_w)):
        o = y[:, i] @ dec_w[i].to(y.dtype).T
        if lay_alphas[i] == 1:
            o = torch.sum(o, dim=-2, keepdim=True)
        out[:, layer_boundaries[i]:layer_boundaries[i + 1]] = o.reshape(B, -1)
    return out


def matrix_enc(x, enc_mat, alpha, r):
    """(B,1,L,N) x (L,E,N) -> (B,1,L,alpha,r)"""
    out = torch.einsum("len,bcln->bcle", enc_mat.to(x.dtype), x)
    return out.reshape(out.shape[0], out.shape[1], out.shape[2], alpha, r)


def matrix_dec(y, dec_mat):
    """(B,1,L,alpha,r) x (L,N,E) -> (B,1,L,N)"""
    flat = y.reshape(y.shape[0], y.shape[1], y.shape[2], -1)
    return torch.einsum("lne,bcle->bcln", dec_mat.to(y.dtype), flat)


def hgcal_matrices(L=28, E=252, N=1988, seed=5):
    """Seeded stand-ins for HGCalConverter's matrices at the probe's density (~0.4 % enc, ~0.6 % dec)."""
    g = torch.Generator().manual_seed(seed)
    enc = torch.rand(L, E, N, generator=g) * (torch.rand(L, E, N, generator=g) < 0.004)
    dec = torch.randn(L, N, E, generator=g) * (torch.rand(L, N, E, generator=g) < 0.006)
    return enc, dec


class GeomEmbedOracle:
    """enc/dec pair handed to DenoiserOracle(geom=...)."""

    def __init__(self, enc, dec):
        self.enc, self.dec = enc, dec

    @classmethod
    def nn_converter(cls, enc_w, dec_w, layer_boundaries, lay_alphas, alpha_out):
        lb = [int(v) for v in layer_boundaries]
        la = [int(v) for v in lay_alphas]
        return cls(lambda x: nn_enc(x, enc_w, lb, la, int(alpha_out)), lambda y: nn_dec(y, dec_w, lb, la, int(alpha_out)))

    @classmethod
    def matrices(cls, enc_mat, dec_mat, alpha, r):
        return cls(lambda x: matrix_enc(x, enc_mat, alpha, r), lambda y: matrix_dec(y, dec_mat))
