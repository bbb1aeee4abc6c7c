"""In-loop geometry embedding: irregular calorimeter voxels <-> the regular (Z, alpha, R) grid of the U-Net.

Host-side mirror of the reference's converters (same class names, constructor arguments, parameter names and
`enc`/`dec` meaning) whose arithmetic runs inside the CUDA library: every converter exports its two linear maps as
CSR matrices (`csr_maps`) which `CaloDiffusion._engine` hands to `cd_unet_set_geom_map`; the device then applies them
inside every denoiser evaluation (prep_input_geom_kernel / geom_decode_kernel).  There is no host compute path: `enc` /
`dec` on these classes go through the library too (`cd_geom_apply`), and fail without a GPU.

  BinningXML      <- utils/XMLHandler.py:11-77 (only what GeomConverter reads: r_edges, n_bin_alpha, bin edges)
  GeomConverter   <- utils/utils.py:659-777
  NNConverter     <- utils/utils.py:575-656   (Dataset 1; trainable per-layer Linear maps `encs.i.weight`, `decs.i.weight`)
  MatrixConverter <- utils/HGCal_utils.py:296-349, 636-663 (Embeder / Decoder of HGCalConverter: per-layer dense
                     matrices `enc_mat (L,E,N)`, `dec_mat (L,N,E)` with boolean masks)
"""
import xml.etree.ElementTree as ET

import numpy as np
import torch
from torch import nn


class BinningXML:
    """The part of the CaloChallenge binning file the geometry converter needs (utils/XMLHandler.py:11-77, 119-149)."""

    def __init__(self, particle_name, filename="binning.xml"):
        root = ET.parse(filename).getroot()
        self.r_edges, self.a_bins = [], []
        found = False
        for particle in root:
            if particle.attrib["name"] == particle_name:
                found = True
                for layer in particle:
                    self.r_edges.append([float(s) for s in layer.attrib.get("r_edges").split(",")])
                    self.a_bins.append(int(layer.attrib.get("n_bin_alpha")))
        if not found:
            raise ValueError(f"Particle {particle_name} not found in {filename}")
        # a layer is relevant when it has at least one radial bin (XMLHandler.SetEtaAndPhiFromPolar)
        self.relevantlayers = [i for i, e in enumerate(self.r_edges) if len(e) > 1]
        self.bin_number = [(len(e) - 1) * a if len(e) > 1 else 0 for e, a in zip(self.r_edges, self.a_bins)]
        self.bin_edges = [0]
        for n in self.bin_number:
            self.bin_edges.append(self.bin_edges[-1] + n)
        # alphaListPerLayer[i][0] has one entry per alpha bin of layer i
        self.alphaListPerLayer = [[[0.0] * a] for a in self.a_bins]

    def GetBinEdges(self):
        return self.bin_edges

    def GetRelevantLayers(self):
        return self.relevantlayers


def write_binning_xml(path, r_edges, a_bins, particle="photon"):
    """Emit a CaloChallenge-style binning file (one <Layer r_edges=... n_bin_alpha=...> per calorimeter layer)."""
    rows = "".join('      <Layer id="%d" r_edges="%s" n_bin_alpha="%d" />\n' % (i, ",".join(repr(float(v)) for v in e), int(a))
                   for i, (e, a) in enumerate(zip(r_edges, a_bins)))
    with open(path, "w") as f:
        f.write('<Bins>\n   <Bin pid="22" etaMin="0" etaMax="130" name="%s">\n%s   </Bin>\n</Bins>\n' % (particle, rows))
    return str(path)


def _csr_from_coo(n_rows, n_cols, rows, cols, vals):
    """Sorted CSR (int32 rowptr/col, fp32 val) from COO triplets."""
    rows = np.asarray(rows, np.int64)
    cols = np.asarray(cols, np.int64)
    vals = np.asarray(vals, np.float32)
    order = np.lexsort((cols, rows))
    rows, cols, vals = rows[order], cols[order], vals[order]
    rowptr = np.zeros(n_rows + 1, np.int64)
    np.add.at(rowptr, rows + 1, 1)
    rowptr = np.cumsum(rowptr)
    return dict(n_rows=int(n_rows), n_cols=int(n_cols), rowptr=rowptr.astype(np.int32), col=cols.astype(np.int32),
                val=np.ascontiguousarray(vals))


class GeomConverter:
    "Convert irregular geometry to regular one (ala CaloChallenge Dataset 1)  — utils/utils.py:659-777"

    def __init__(self, bins=None, all_r_edges=None, lay_r_edges=None, alpha_out=1, lay_alphas=None):
        self.layer_boundaries = []
        self.bins = None
        if bins is not None:
            self.layer_boundaries = np.unique(bins.GetBinEdges())
            rel_layers = bins.GetRelevantLayers()
            lay_alphas = [len(bins.alphaListPerLayer[idx][0]) for idx, redge in enumerate(bins.r_edges) if len(redge) > 1]
            alpha_out = int(np.amax(lay_alphas))
            lay_r_edges = [bins.r_edges[l] for l in rel_layers]
            all_r_edges = torch.unique(torch.tensor([e for lay in lay_r_edges for e in lay], dtype=torch.float32))
        self.all_r_edges = all_r_edges
        self.lay_r_edges = lay_r_edges
        self.alpha_out = alpha_out
        self.lay_alphas = lay_alphas
        self.num_layers = len(self.lay_r_edges)
        self.all_r_areas = all_r_edges[1:] ** 2 - all_r_edges[:-1] ** 2
        self.dim_r_out = len(all_r_edges) - 1
        self.weight_mats = []
        edges = [float(v) for v in all_r_edges]
        for lay in lay_r_edges:
            dim_in = len(lay) - 1
            w = torch.zeros((self.dim_r_out, dim_in))
            for ir in range(dim_in):
                lo, hi = edges.index(float(np.float32(lay[ir]))), edges.index(float(np.float32(lay[ir + 1])))
                orig_area = lay[ir + 1] ** 2 - lay[ir] ** 2
                w[lo:hi, ir] = self.all_r_areas[lo:hi] / orig_area   # split proportional to bin area
            self.weight_mats.append(w)

    def flat_size(self):
        return int(self.layer_boundaries[-1])

    def grid_shape(self):
        return (self.num_layers, int(self.alpha_out), int(self.dim_r_out))

    def layer_maps_to_csr(self, enc_w, dec_w):
        """CSR forms of `enc` (grid <- flat) and `dec` (flat <- grid) for per-layer matrices enc_w[i] (R_out, r_i) and
        dec_w[i] (r_i, R_out), including the alpha handling of utils/utils.py:617-630 / :641-650 (a layer with one alpha
        bin is spread evenly over alpha_out on the way in and summed over alpha on the way out)."""
        A, Ro = int(self.alpha_out), int(self.dim_r_out)
        n_grid, n_flat = self.num_layers * A * Ro, self.flat_size()
        er, ec, ev, dr, dc, dv = [], [], [], [], [], []
        for i in range(self.num_layers):
            a_i = int(self.lay_alphas[i]) if self.lay_alphas is not None else A
            if a_i not in (1, A):
                raise ValueError("Num alpha bins for layer %i is %i. Don't know how to handle" % (i, a_i))
            We = np.asarray(enc_w[i], np.float32)
            Wd = np.asarray(dec_w[i], np.float32)
            r_i = We.shape[1]
            lb = int(self.layer_boundaries[i])
            if int(self.layer_boundaries[i + 1]) - lb != a_i * r_i:
                raise ValueError(f"layer {i}: {a_i}x{r_i} bins do not fill [{lb},{self.layer_boundaries[i + 1]})")
            rr, jj = np.meshgrid(np.arange(Ro), np.arange(r_i), indexing="ij")
            for a in range(A):
                a_src = 0 if a_i == 1 else a
                cell = (i * A + a) * Ro + rr
                flat = lb + a_src * r_i + jj
                er.append(cell.ravel()); ec.append(flat.ravel())
                ev.append((We / np.float32(A) if a_i == 1 else We).ravel())
                dr.append(flat.T.ravel()); dc.append(cell.T.ravel()); dv.append(Wd.ravel())
        cat = np.concatenate
        return (_csr_from_coo(n_grid, n_flat, cat(er), cat(ec), cat(ev)),
                _csr_from_coo(n_flat, n_grid, cat(dr), cat(dc), cat(dv)))


class _DeviceMaps:
    """Shared enc/dec plumbing: the CSR maps live in int32/fp32 device tensors and are applied by the CUDA library."""

    def _device_csr(self, device):
        key = (str(device), self._weights_version())
        if getattr(self, "_csr_cache_key", None) != key:
            enc, dec = self.csr_maps()
            to = lambda a: torch.from_numpy(a).to(device)
            self._csr_dev = [{k: (to(v) if isinstance(v, np.ndarray) else v) for k, v in m.items()} for m in (enc, dec)]
            self._csr_cache_key = key
        return self._csr_dev

    def _apply_map(self, which, x):
        from . import _lib
        lib = _lib.load()
        if not x.is_cuda:
            raise RuntimeError("geometry enc/dec run in the CUDA library: pass a CUDA tensor (there is no CPU path)")
        m = self._device_csr(x.device)[which]
        B = x.shape[0]
        xin = x.to(torch.float32).reshape(B, m["n_cols"]).contiguous()
        out = torch.empty((B, m["n_rows"]), device=x.device, dtype=torch.float32)
        _lib.check(lib.cd_geom_apply(m["rowptr"].data_ptr(), m["col"].data_ptr(), m["val"].data_ptr(), m["n_rows"], m["n_cols"],
                                     xin.data_ptr(), out.data_ptr(), B, torch.cuda.current_stream().cuda_stream), "geom_apply")
        return out


class NNConverter(nn.Module, _DeviceMaps):
    "Convert irregular geometry to regular one, initialized with regular geometric conversion, but uses trainable linear map"

    def __init__(self, geomconverter=None, bins=None, hidden_size=32):
        super().__init__()
        self.gc = geomconverter if geomconverter is not None else GeomConverter(bins)
        self.encs = nn.ModuleList([])
        self.decs = nn.ModuleList([])
        eps = 1e-5
        for w in self.gc.weight_mats:   # utils/utils.py:588-608: area-weighted map + 1e-5 noise; decoder = pseudo-inverse
            lay = nn.Linear(w.shape[1], self.gc.dim_r_out, bias=False)
            lay.weight.data = w + eps * torch.randn_like(w)
            self.encs.append(lay)
            inv = torch.linalg.pinv(w)
            inv_lay = nn.Linear(self.gc.dim_r_out, w.shape[1], bias=False)
            inv_lay.weight.data = inv + eps * torch.randn_like(inv)
            self.decs.append(inv_lay)

    def _weights_version(self):
        return tuple(p._version for p in self.parameters()) + tuple(p.data_ptr() for p in self.parameters())

    def csr_maps(self):
        return self.gc.layer_maps_to_csr([l.weight.detach().cpu().numpy() for l in self.encs],
                                         [l.weight.detach().cpu().numpy() for l in self.decs])

    def flat_shape(self):
        return (self.gc.flat_size(),)

    def enc(self, x):
        """(B, n_flat) -> (B, 1, layers, alpha_out, dim_r_out)   utils/utils.py:610-634"""
        return self._apply_map(0, x).reshape((x.shape[0], 1) + self.gc.grid_shape())

    def dec(self, x):
        """(B, 1, layers, alpha_out, dim_r_out) -> (B, n_flat)   utils/utils.py:636-653"""
        return self._apply_map(1, x)

    def forward(self, x):
        return self.enc(x)


class MatrixConverter(nn.Module, _DeviceMaps):
    """Per-layer dense embedding matrices of the HGCal converter (Embeder / Decoder, utils/HGCal_utils.py:296-349):
    enc_mat (L, A*R, N), dec_mat (L, N, A*R), optional boolean masks of the same shapes (applied when `trainable`,
    as the reference does).  Building the matrices from the HGCal geometry pickle (HGCalConverter.init, init_map) is the
    reference's host code and stays there: pass its `enc_mat`/`dec_mat` (or a checkpoint's) in."""

    def __init__(self, bins, enc_mat, dec_mat, enc_mask=None, dec_mask=None, trainable=False):
        super().__init__()
        self.num_layers, self.num_alpha_bins, self.num_r_bins = int(bins[-3]), int(bins[-2]), int(bins[-1])
        enc_mat = torch.as_tensor(enc_mat, dtype=torch.float32)
        dec_mat = torch.as_tensor(dec_mat, dtype=torch.float32)
        E = self.num_alpha_bins * self.num_r_bins
        if enc_mat.shape[0] != self.num_layers or enc_mat.shape[1] != E or tuple(dec_mat.shape) != (enc_mat.shape[0], enc_mat.shape[2], E):
            raise ValueError(f"embedding matrices {tuple(enc_mat.shape)} / {tuple(dec_mat.shape)} do not match bins {tuple(bins)}")
        self.max_ncell = int(enc_mat.shape[2])
        self.trainable = trainable
        if trainable:
            self.enc_mat, self.dec_mat = nn.Parameter(enc_mat), nn.Parameter(dec_mat)
        else:
            self.register_buffer("enc_mat", enc_mat)
            self.register_buffer("dec_mat", dec_mat)
        self.register_buffer("enc_mask", torch.ones_like(enc_mat, dtype=torch.bool) if enc_mask is None else torch.as_tensor(enc_mask, dtype=torch.bool))
        self.register_buffer("dec_mask", torch.ones_like(dec_mat, dtype=torch.bool) if dec_mask is None else torch.as_tensor(dec_mask, dtype=torch.bool))

    def _weights_version(self):
        return (self.enc_mat._version, self.dec_mat._version, self.enc_mat.data_ptr(), self.dec_mat.data_ptr())

    def csr_maps(self):
        L, E, N = self.enc_mat.shape
        out = []
        for mat, mask, (nr, nc) in ((self.enc_mat, self.enc_mask, (E, N)), (self.dec_mat, self.dec_mask, (N, E))):
            m = mat.detach().cpu().numpy()
            if self.trainable:
                m = m * mask.cpu().numpy()
            l, r, c = np.nonzero(m)
            out.append(_csr_from_coo(L * nr, L * nc, l * nr + r, l * nc + c, m[l, r, c]))
        return tuple(out)

    def flat_shape(self):
        return (1, self.num_layers, self.max_ncell)

    def enc(self, x):
        """(B, 1, L, N) -> (B, 1, L, A, R)   HGCal_utils.py:315-320, 636-640"""
        return self._apply_map(0, x).reshape(x.shape[0], 1, self.num_layers, self.num_alpha_bins, self.num_r_bins)

    def dec(self, x, sparse_decoding=False, sparse_per_batch=False):
        """(B, 1, L, A, R) -> (B, 1, L, N)   HGCal_utils.py:340-349, 659-663 (averaging decode only)"""
        if sparse_decoding:
            raise ValueError("sparse (sampled) decoding is not on the B200 path")
        return self._apply_map(1, x).reshape(x.shape[0], 1, self.num_layers, self.max_ncell)

    def forward(self, x):
        return self.enc(x)
