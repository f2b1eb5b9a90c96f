"""Channel-axis transforms either side of the HGA path, on the GPU.

The reference applies them as numpy callables in `pre_transforms` / `post_transforms` of `HighGammaExtractor`
(/root/reference/local/units.py:147-148,158-159):

  * column gathers        `SelectElectrodesFromBothGrids`, `SelectElectrodesOverSpeechAreas`   local/common.py:16-58
  * per-grid common average referencing   `CommonAverageReferencing`                           local/common.py:308-343
  * bad-channel patching  `BadChannelCorrection`                                               local/common.py:220-284

Every one of them, and every chain of them that contains at most one averaging stage, is the row-wise map

    out[t, j] = alpha[j] * x[t, sel[j]] + beta[j] * mean_{c in group grp[j]} x[t, c]

which `hga_chanmap_*` (include/hga_b200.h, csrc/hga_kernels_pre.cuh) runs as one HBM-bound kernel.  `ChannelMapSpec`
holds that map in RAW input columns; `fold(transforms, c_in)` recognises the reference's own objects by their attributes
(so `from local.common import ...` keeps working unchanged) as well as the classes below, and composes a chain into one
spec.  Chains that cannot be folded (two averaging stages, unknown callables) return None: the extractor then calls the
user's callables exactly as the reference does.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass, field
from typing import Callable, List, Optional, Sequence

import numpy as np

from . import _lib
from ._lib import check


MAX_GROUPS = 64   # kChanMapMaxGroups (csrc/hga_kernels_pre.cuh)


@dataclass
class ChannelMapSpec:
    c_in: int
    sel: np.ndarray                      # (c_out,) int32 raw column of the direct term, -1 = none
    grp: np.ndarray                      # (c_out,) int32 group of the mean term, -1 = none
    alpha: np.ndarray                    # (c_out,) float64
    beta: np.ndarray                     # (c_out,) float64
    groups: List[np.ndarray] = field(default_factory=list)   # raw columns of each group

    @property
    def c_out(self) -> int:
        return int(len(self.sel))

    @property
    def has_mean(self) -> bool:
        return bool((self.grp >= 0).any())

    @staticmethod
    def identity(c_in: int) -> "ChannelMapSpec":
        return ChannelMapSpec(c_in, np.arange(c_in, dtype=np.int32), np.full(c_in, -1, np.int32), np.ones(c_in), np.zeros(c_in), [])

    def then_gather(self, idx: Sequence[int]) -> "ChannelMapSpec":
        """self followed by `data[:, idx]`."""
        idx = np.asarray(idx, dtype=np.int64).reshape(-1)
        if idx.size and (idx.min() < -self.c_out or idx.max() >= self.c_out):
            raise IndexError(f"index {idx.max()} is out of bounds for axis 1 with size {self.c_out}")
        idx = np.where(idx < 0, idx + self.c_out, idx)
        return ChannelMapSpec(self.c_in, self.sel[idx].copy(), self.grp[idx].copy(), self.alpha[idx].copy(), self.beta[idx].copy(),
                              [g.copy() for g in self.groups])

    def then_means(self, targets: Sequence[np.ndarray], sources: Sequence[np.ndarray], alpha: float, beta: float) -> Optional["ChannelMapSpec"]:
        """self followed by: for each k, columns targets[k] <- alpha * column + beta * mean(columns sources[k]).
        Only foldable when self is a pure gather (no mean term yet, unit weights)."""
        if self.has_mean or not np.all(self.alpha == 1.0) or np.any(self.sel < 0):
            return None
        out = ChannelMapSpec(self.c_in, self.sel.copy(), self.grp.copy(), self.alpha.copy(), self.beta.copy(), [])
        for tgt, src in zip(targets, sources):
            tgt, src = np.asarray(tgt, dtype=np.int64).reshape(-1), np.asarray(src, dtype=np.int64).reshape(-1)
            if src.size == 0:
                return None  # numpy would produce NaN (mean of an empty slice); leave that to the reference's own code
            g = len(out.groups)
            out.groups.append(self.sel[src].astype(np.int32))
            out.grp[tgt] = g
            out.alpha[tgt] = alpha
            out.beta[tgt] = beta
            if alpha == 0.0:
                out.sel[tgt] = -1       # no direct term: a non-finite value in the replaced column must not leak in
        return out


# ---------------------------------------------------------------------------------------------- recognising transforms
def _cname(t) -> str:
    return type(t).__name__


def _fold_one(spec: ChannelMapSpec, t) -> Optional[ChannelMapSpec]:
    if isinstance(t, ChannelSelection):
        return spec.then_gather(t.mapping)
    if isinstance(t, CommonAverageReferencing) or (_cname(t) == "CommonAverageReferencing" and hasattr(t, "selection_masks_computation")
                                                   and hasattr(t, "selection_masks_application")):
        appl = [np.flatnonzero(np.asarray(m)) for m in t.selection_masks_application]
        comp = [np.flatnonzero(np.asarray(m)) for m in t.selection_masks_computation]
        if any(len(np.asarray(m)) != spec.c_out for m in t.selection_masks_application):
            return None
        # later grids see the ORIGINAL data for their means (local/common.py:338-341 reads `data`, writes `result`), and
        # a column belonging to two grids would be referenced twice: only disjoint grids fold into one map
        if len(appl) > 1 and len(np.unique(np.concatenate(appl))) != sum(len(a) for a in appl):
            return None
        return spec.then_means(appl, comp, 1.0, -1.0)
    if isinstance(t, BadChannelCorrection) or (_cname(t) == "BadChannelCorrection" and hasattr(t, "patches")):
        tg, sr = [], []
        for loc, neigh in t.patches:
            loc, neigh = np.asarray(loc).reshape(-1), np.asarray(neigh).reshape(-1)
            if loc.size == 0:
                continue  # the bad channel is not in the layout: numpy assigns to an empty selection
            tg.append(loc)
            sr.append(neigh)
        if any(np.intersect1d(a, b).size for i, a in enumerate(tg) for b in tg[i + 1:]):
            return None
        return spec.then_means(tg, sr, 0.0, 1.0) if tg else spec
    for attr in ("grid_mapping", "speech_grid_mapping"):   # the reference's two selection classes (local/common.py:16-58)
        if _cname(t).startswith("SelectElectrodes") and hasattr(t, attr):
            return spec.then_gather(getattr(t, attr))
    return None


def fold(transforms: Optional[Sequence[Callable]], c_in: int) -> Optional[ChannelMapSpec]:
    """One spec for the whole chain applied to (T, c_in) data, or None if it cannot be expressed as one map."""
    if not transforms:
        return None
    spec = ChannelMapSpec.identity(int(c_in))
    try:
        for t in transforms:
            spec = _fold_one(spec, t)
            if spec is None:
                return None
    except (IndexError, ValueError, TypeError):
        return None
    if len(spec.groups) > MAX_GROUPS:      # more averaging groups than the kernel's table holds: the callables run as they are
        return None
    return spec


# ---------------------------------------------------------------------------------------------- device object
class _RawDtypes(dict):
    """torch dtype -> HGA_RAW_* code of include/hga_b200.h (built on first use: torch is imported lazily here)."""

    def _fill(self):
        if not self:
            import torch
            self.update({torch.float64: 0, torch.float32: 1, torch.int32: 2, torch.int16: 3})

    def __contains__(self, k):
        self._fill()
        return dict.__contains__(self, k)

    def __getitem__(self, k):
        self._fill()
        return dict.__getitem__(self, k)


RAW_DTYPES = _RawDtypes()
RAW_NUMPY = {np.dtype(np.float64): 0, np.dtype(np.float32): 1, np.dtype(np.int32): 2, np.dtype(np.int16): 3}


class ChannelMap:
    """A folded chain on one device: `apply(x)` for (T, c_in) float64 numpy arrays or torch CUDA tensors."""

    def __init__(self, spec: ChannelMapSpec, device: Optional[int] = None):
        import torch

        self._h = None
        lib = _lib.load()
        _lib.require_device()
        self.spec = spec
        self.device = int(torch.cuda.current_device() if device is None else device)
        sel = np.ascontiguousarray(spec.sel, dtype=np.int32)
        grp = np.ascontiguousarray(spec.grp, dtype=np.int32)
        alpha = np.ascontiguousarray(spec.alpha, dtype=np.float64)
        beta = np.ascontiguousarray(spec.beta, dtype=np.float64)
        gptr = np.zeros(len(spec.groups) + 1, dtype=np.int32)
        gptr[1:] = np.cumsum([len(g) for g in spec.groups])
        gidx = np.ascontiguousarray(np.concatenate(spec.groups) if spec.groups else np.zeros(0), dtype=np.int32)
        p32 = lambda a: a.ctypes.data_as(C.POINTER(C.c_int32))
        h = C.c_void_p()
        check(lib.hga_chanmap_create(self.device, spec.c_in, spec.c_out, p32(sel), p32(grp), alpha.ctypes.data, beta.ctypes.data,
                                     len(spec.groups), p32(gptr), p32(gidx), C.byref(h)))
        self._h = h

    def __del__(self):
        if getattr(self, "_h", None) is not None:
            try:
                _lib.load().hga_chanmap_destroy(self._h)
            except Exception:
                pass
            self._h = None

    def apply(self, x, out=None):
        import torch

        lib = _lib.load()
        if isinstance(x, torch.Tensor):
            if not x.is_cuda or x.dtype != torch.float64 or x.dim() != 2 or x.shape[1] != self.spec.c_in:
                raise ValueError(f"expected a CUDA float64 (T, {self.spec.c_in}) tensor")
            if x.stride(1) != 1:
                x = x.contiguous()
            T = x.shape[0]
            if out is None:
                out = torch.empty((T, self.spec.c_out), dtype=torch.float64, device=x.device)
            with torch.cuda.device(self.device):
                check(lib.hga_chanmap_apply(self._h, x.data_ptr(), T, max(x.stride(0), self.spec.c_in), out.data_ptr(), out.stride(0),
                                            int(torch.cuda.current_stream(self.device).cuda_stream)))
            return out
        x = np.asarray(x)
        if x.ndim != 2 or x.shape[1] != self.spec.c_in:
            raise IndexError(f"expected (T, {self.spec.c_in}) data, got {x.shape}")
        x = np.ascontiguousarray(x, dtype=np.float64)
        res = np.empty((x.shape[0], self.spec.c_out), dtype=np.float64) if out is None else out
        check(lib.hga_chanmap_apply_host(self._h, x.ctypes.data, x.shape[0], x.shape[1], res.ctypes.data, res.shape[1]))
        return res

    def apply_raw(self, x, gain=None, out=None):
        """The map over raw amplifier counts: a CUDA (T, c_in) tensor of float64 / float32 / int32 / int16, each column
        scaled by `gain` (scalar or (c_in,), float64) at the load -- `BCI2000MatFile.signals()` = `signal * gain`
        (local/common.py:456-459) fused in front of the chain.  Returns a float64 (T, c_out) CUDA tensor."""
        import torch

        lib = _lib.load()
        if not isinstance(x, torch.Tensor) or not x.is_cuda or x.dim() != 2 or x.shape[1] != self.spec.c_in:
            raise ValueError(f"expected a CUDA (T, {self.spec.c_in}) tensor")
        if x.dtype not in RAW_DTYPES:
            raise TypeError(f"raw element type {x.dtype} is not one of float64, float32, int32, int16")
        if x.stride(1) != 1:
            x = x.contiguous()
        g = None
        if gain is not None:
            g = np.broadcast_to(np.asarray(gain, dtype=np.float64).reshape(-1), (self.spec.c_in,))
            g = torch.from_numpy(np.array(g)).to(x.device)
        T = x.shape[0]
        if out is None:
            out = torch.empty((T, self.spec.c_out), dtype=torch.float64, device=x.device)
        with torch.cuda.device(self.device):
            check(lib.hga_chanmap_apply_raw(self._h, x.data_ptr(), RAW_DTYPES[x.dtype], None if g is None else g.data_ptr(), T,
                                            max(x.stride(0), self.spec.c_in), out.data_ptr(), out.stride(0),
                                            int(torch.cuda.current_stream(self.device).cuda_stream)))
        return out


class _MapBacked:
    """Callable transform whose arithmetic runs in the channel-map kernel (built lazily, per input width)."""

    def _spec_for(self, c_in: int) -> ChannelMapSpec:
        spec = _fold_one(ChannelMapSpec.identity(c_in), self)
        if spec is None:
            raise ValueError(f"{_cname(self)} cannot be applied to {c_in}-column data")
        return spec

    def __call__(self, data):
        c_in = int(data.shape[1])
        cache = self.__dict__.setdefault("_maps", {})
        if c_in not in cache:
            cache[c_in] = ChannelMap(self._spec_for(c_in))
        return cache[c_in].apply(data)


# ---------------------------------------------------------------------------------------------- mirrored classes
class ChannelSelection(_MapBacked):
    """`data[:, mapping]` — the generic form of the reference's two montage-specific selection classes
    (local/common.py:16-58; their index tables are subject data and stay in the reference)."""

    def __init__(self, mapping: Sequence[int]):
        self.mapping = np.asarray(mapping, dtype=np.int64).reshape(-1)

    def __len__(self) -> int:
        return len(self.mapping)

    def __repr__(self) -> str:
        return f"Channels: {', '.join(map(str, self.mapping + 1))}"


class CommonAverageReferencing(_MapBacked):
    """local/common.py:308-365.  Per electrode grid, subtract at every time sample the mean over the grid's channels
    that are not excluded.  Data is (T, E); `layout[i]` is the channel number held by column i."""

    def __init__(self, exclude_channels: List[int], grids: List[np.ndarray], layout: np.ndarray):
        self.grids = grids
        self.layout = np.asarray(layout)
        excluded = set(int(c) for c in exclude_channels)
        self.selection_masks_application = []
        self.selection_masks_computation = []
        for grid in grids:
            members = set(int(c) for c in np.asarray(grid).reshape(-1))
            appl = np.array([int(c) in members for c in self.layout], dtype=bool)
            comp = appl.copy()
            for ch in excluded & members:
                hit = np.flatnonzero(self.layout == ch)
                # the reference clears column argmax(layout == ch): the first match, or column 0 when there is none
                comp[hit[0] if hit.size else 0] = False
            self.selection_masks_application.append(appl)
            self.selection_masks_computation.append(comp)

    def __repr__(self) -> str:
        lines = [f"CommonAverageReferencing: {len(self.grids)} grids"]
        for i, (a, c) in enumerate(zip(self.selection_masks_application, self.selection_masks_computation)):
            lines.append(f"Grid {i}: mean over {int(c.sum())} of {int(a.sum())} channels")
        return "\n".join(lines)


class BadChannelCorrection(_MapBacked):
    """local/common.py:220-305.  Replace every bad channel by the mean of its non-bad 8-neighbours on its grid."""

    def __init__(self, bad_channels: List[int], grids: List[np.ndarray], layout: np.ndarray):
        self.grids = [np.asarray(g) for g in grids]
        self.layout = np.asarray(layout)
        bad = [int(c) for c in bad_channels]
        self.masks = [~np.isin(g, bad) for g in self.grids]
        self.footprint = np.ones((3, 3), dtype=bool)
        self.footprint[1, 1] = False
        self.patches = []
        for ch in bad:
            gi = next((i for i, g in enumerate(self.grids) if (g == ch).any()), None)
            if gi is None:
                raise IndexError('Channel could not be found in given grids.')
            grid, ok = self.grids[gi], self.masks[gi]
            near = np.zeros(grid.shape, dtype=bool)
            for r, c in zip(*np.nonzero(grid == ch)):
                r0, r1, c0, c1 = max(r - 1, 0), min(r + 2, grid.shape[0]), max(c - 1, 0), min(c + 2, grid.shape[1])
                near[r0:r1, c0:c1] = True
            neighbours = grid[near & ok]          # row-major order, the bad channel itself is masked out by `ok`
            cols = [np.flatnonzero(self.layout == n) for n in neighbours]
            self.patches.append((np.flatnonzero(self.layout == ch), np.concatenate(cols) if cols else np.zeros(0, dtype=np.int64)))

    def __len__(self) -> int:
        return len(self.patches)

    def __repr__(self) -> str:
        items = [f"{self.layout[loc].item() if loc.size == 1 else list(self.layout[loc])} -> {[int(self.layout[i]) for i in nb]}"
                 for loc, nb in self.patches]
        return f"Correcting {len(self.patches)} bad channels: {', '.join(items)}"
