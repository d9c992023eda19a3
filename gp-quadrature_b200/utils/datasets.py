"""Dataset ingestion next to the hot path: what the reference's loaders do AFTER the NetCDF
file has been parsed (oisst_experiment/load_oisst.py:93-158, era5_experiment/load_era5.py:16-69) -- gridded field ->
(x, y) point lists, longitude convention, Kelvin -> Celsius, uniform subsampling -- plus the notebooks' normalisation to
the unit box and a one-time, h-independent SPATIAL PRESORT that is persisted beside the data.

NetCDF parsing itself needs h5py / xarray, which this image does not have ("unavailable on this box"):
`load_points` reads the same content from .pt / .npy / .npz files.  Everything here is plain torch and runs on
whatever device the tensors live on (the O(N) passes over 1e8 points belong on the GPU).
"""
import os

import numpy as np
import torch


def load_points(path, device=None, dtype=None):
    """(x (N, d), y (N,)) from a .pt file holding [x, y] / (x, y) / {'x':, 'y':} (the reference's data/*.pt fixtures),
    a .npz with arrays x, y, or a pair of .npy files `path` (x) and `path` with _y before the extension."""
    ext = os.path.splitext(path)[1].lower()
    if ext == ".pt":
        obj = torch.load(path, map_location="cpu")
        x, y = (obj["x"], obj["y"]) if isinstance(obj, dict) else (obj[0], obj[1])
    elif ext == ".npz":
        z = np.load(path)
        x, y = torch.from_numpy(z["x"]), torch.from_numpy(z["y"])
    elif ext == ".npy":
        x = torch.from_numpy(np.load(path))
        y = torch.from_numpy(np.load(path[:-4] + "_y.npy"))
    else:
        raise ValueError(f"unsupported point file {path!r}: use .pt, .npz or .npy")
    x, y = torch.as_tensor(x), torch.as_tensor(y)
    if x.ndim == 1:
        x = x[:, None]
    if dtype is not None:
        x, y = x.to(dtype), y.to(dtype)
    if device is not None:
        x, y = x.to(device), y.to(device)
    return x.contiguous(), y.reshape(-1).contiguous()


def grid_to_points(lon, lat, values, lon_range="-180_180", n_sub=None, seed=0, kelvin_to_celsius=False):
    """A field on a (n_lat, n_lon) lon/lat grid -> valid pixels as points, as load_oisst (load_oisst.py:116-146) and
    load_era5 (load_era5.py:39-66) do: longitudes wrapped to [-180, 180) (and re-ordered) or kept in [0, 360), NaN /
    masked pixels dropped, optional uniform subsample drawn with numpy's default_rng(seed).choice exactly like the
    reference, x = (longitude, latitude) in degrees (float64), y = the values."""
    lon = torch.as_tensor(lon, dtype=torch.float64)
    lat = torch.as_tensor(lat, dtype=torch.float64)
    values = torch.as_tensor(values)
    if values.ndim == 3:                       # singleton time axis (load_era5.py:44-45)
        values = values[0]
    if lon_range == "-180_180":
        lon = torch.where(lon >= 180, lon - 360, lon)
        order = torch.argsort(lon)
        lon, values = lon[order], values[:, order.to(values.device)]
    elif lon_range != "0_360":
        raise ValueError("lon_range must be '-180_180' or '0_360'")
    LON, LAT = torch.meshgrid(lon.to(values.device), lat.to(values.device), indexing="xy")
    valid = torch.isfinite(values)
    x = torch.stack([LON[valid], LAT[valid]], dim=1)
    y = values[valid].to(torch.float64)
    if kelvin_to_celsius:
        y = y - 273.15
    if n_sub is not None and n_sub < y.numel():
        inds = torch.from_numpy(np.random.default_rng(seed).choice(y.numel(), size=n_sub, replace=False)).to(x.device)
        x, y = x[inds], y[inds]
    return x.contiguous(), y.contiguous()


def normalize_unit_box(x):
    """The notebooks' normalisation before EFGPND: shift to the origin and divide by the LARGEST axis extent, so the
    points lie in [0, 1]^d with the aspect ratio kept.  Returns (x_unit, offset (d,), scale)."""
    lo, hi = torch.aminmax(x, dim=0)
    scale = (hi - lo).max()
    scale = scale if float(scale) > 0 else torch.ones_like(scale)
    return (x - lo) / scale, lo, scale


def _spread_bits(v, nbits, d):
    """Insert d-1 zero bits between the low `nbits` bits of v (int64)."""
    out = torch.zeros_like(v)
    for b in range(nbits):
        out |= ((v >> b) & 1) << (b * d)
    return out


def spatial_presort(x, y=None, bits=None):
    """One-time, h-INDEPENDENT spatial ordering of the points: Morton (Z-order) key of the point's cell on a 2^bits
    grid per axis over the bounding box.  Any later fine grid (every hyperparameter step changes h and the grid size)
    sees spatially coherent input: the level-1 bucket scatter of the library's cell sort and both passes of the
    un-permute then move long contiguous runs instead of 2048 / 128-point fragments.  Returns (x_sorted, y_sorted, perm)
    with x_sorted = x[perm]."""
    N, d = x.shape
    if bits is None:
        bits = max(1, min(20, 62 // d))
    lo, hi = torch.aminmax(x, dim=0)
    ext = torch.where(hi > lo, hi - lo, torch.ones_like(hi))
    cells = ((x - lo) / ext * (2 ** bits)).to(torch.int64).clamp_(0, 2 ** bits - 1)
    key = torch.zeros(N, dtype=torch.int64, device=x.device)
    for a in range(d):
        key |= _spread_bits(cells[:, a], bits, d) << a
    perm = torch.argsort(key)
    return x[perm].contiguous(), (None if y is None else y[perm].contiguous()), perm


def save_presorted(path, x, y, perm=None, meta=None):
    """Persist the presorted points beside the data (a .pt file: x, y, the permutation back to file order, metadata)."""
    torch.save({"x": x.cpu(), "y": y.cpu(), "perm": None if perm is None else perm.cpu(), "meta": meta or {},
                "presorted": True}, path)


def load_presorted(path, device=None):
    obj = torch.load(path, map_location="cpu")
    if not (isinstance(obj, dict) and obj.get("presorted")):
        raise ValueError(f"{path!r} is not a presorted point file (see save_presorted)")
    x, y, perm = obj["x"], obj["y"], obj.get("perm")
    if device is not None:
        x, y = x.to(device), y.to(device)
        perm = None if perm is None else perm.to(device)
    return x, y, perm, obj.get("meta", {})
