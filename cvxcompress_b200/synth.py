"""Deterministic synthetic volumes.  numpy versions are canonical (used by the
parity tests); the torch versions evaluate the same formulas on the device for bench.py.

G1  golden sinusoid, Test_With_Generated_Input.cpp:45-51: v = (float)sin(iz*pi/320*10), constant in x,y
G2  seismic-like, after Read_Raw_Volume.cpp:18-39: v = (float)(sin(r/10) + u/100), r = distance from the
    centre cell, u in [0,1) from a counter-based hash of the linear cell index (seed 20241027)
G3  block fill, after Fill_Block CvxCompress.cpp:577-584: U[0,1) per cell from the same hash
Volumes are indexed [z, y, x] (x fastest), matching CvxCompress.hxx:27.
"""
import numpy as np

SEED = 20241027
_M32 = 0xFFFFFFFF


def _hash32_np(idx, seed=SEED):
    """lowbias32-style integer hash on uint64 lanes holding 32-bit values (exact in numpy and torch)."""
    x = (idx.astype(np.uint64) ^ np.uint64(seed)) & np.uint64(_M32)
    x ^= (idx.astype(np.uint64) >> np.uint64(32)) * np.uint64(0x9E3779B1) & np.uint64(_M32)
    x ^= x >> np.uint64(16)
    x = (x * np.uint64(0x7FEB352D)) & np.uint64(_M32)
    x ^= x >> np.uint64(15)
    x = (x * np.uint64(0x846CA68B)) & np.uint64(_M32)
    x ^= x >> np.uint64(16)
    return x


def uniform01_np(idx, seed=SEED):
    return ((_hash32_np(idx, seed) >> np.uint64(8)).astype(np.float64) / float(1 << 24))


def g1(nx, ny, nz):
    z = np.arange(nz, dtype=np.float64)
    col = np.sin(z * np.pi / 320.0 * 10.0).astype(np.float32)
    return np.ascontiguousarray(np.broadcast_to(col[:, None, None], (nz, ny, nx)))


def g2(nx, ny, nz, seed=SEED):
    z = np.arange(nz, dtype=np.float64)[:, None, None] - (nz - 1) / 2.0
    y = np.arange(ny, dtype=np.float64)[None, :, None] - (ny - 1) / 2.0
    x = np.arange(nx, dtype=np.float64)[None, None, :] - (nx - 1) / 2.0
    r = np.sqrt(x * x + y * y + z * z)
    idx = np.arange(nx * ny * nz, dtype=np.uint64).reshape(nz, ny, nx)
    return (np.sin(r / 10.0) + uniform01_np(idx, seed) / 100.0).astype(np.float32)


def g2_slab(nx, ny, nz, z0=0, nz_total=None, seed=SEED):
    """z-slab [z0, z0+nz) of the nz_total-deep G2 volume (same cells as g2(nx, ny, nz_total)[z0:z0+nz])."""
    nzt = nz if nz_total is None else nz_total
    z = np.arange(z0, z0 + nz, dtype=np.float64)[:, None, None] - (nzt - 1) / 2.0
    y = np.arange(ny, dtype=np.float64)[None, :, None] - (ny - 1) / 2.0
    x = np.arange(nx, dtype=np.float64)[None, None, :] - (nx - 1) / 2.0
    r = np.sqrt(x * x + y * y + z * z)
    idx = (np.arange(nx * ny * nz, dtype=np.uint64) + np.uint64(z0 * nx * ny)).reshape(nz, ny, nx)
    return (np.sin(r / 10.0) + uniform01_np(idx, seed) / 100.0).astype(np.float32)


def g3(n, seed=SEED, lo=0.0, hi=1.0):
    idx = np.arange(n, dtype=np.uint64)
    return (lo + (hi - lo) * uniform01_np(idx, seed)).astype(np.float32)


# ---- device-side generators (bench only) -----------------------------------------------------
def _hash32_torch(idx, seed=SEED):
    x = (idx ^ seed) & _M32
    x = x ^ (((idx >> 32) * 0x9E3779B1) & _M32)
    x = x ^ (x >> 16)
    x = (x * 0x7FEB352D) & _M32
    x = x ^ (x >> 15)
    x = (x * 0x846CA68B) & _M32
    x = x ^ (x >> 16)
    return x


def g1_torch(nx, ny, nz, device, z0=0, nz_total=None):
    import torch
    z = torch.arange(z0, z0 + nz, dtype=torch.float64, device=device)
    col = torch.sin(z * np.pi / 320.0 * 10.0).to(torch.float32)
    return col[:, None, None].expand(nz, ny, nx).contiguous()


def g2_torch(nx, ny, nz, device, z0=0, nz_total=None, seed=SEED):
    """z-slab [z0, z0+nz) of the nz_total-deep G2 volume."""
    import torch
    nzt = nz if nz_total is None else nz_total
    out = torch.empty((nz, ny, nx), dtype=torch.float32, device=device)
    y = torch.arange(ny, dtype=torch.float64, device=device)[:, None] - (ny - 1) / 2.0
    x = torch.arange(nx, dtype=torch.float64, device=device)[None, :] - (nx - 1) / 2.0
    rxy = x * x + y * y
    plane = torch.arange(nx * ny, dtype=torch.int64, device=device).reshape(ny, nx)
    for k in range(nz):
        zz = (z0 + k) - (nzt - 1) / 2.0
        r = torch.sqrt(rxy + zz * zz)
        idx = plane + (z0 + k) * nx * ny
        u = (_hash32_torch(idx, seed) >> 8).to(torch.float64) / float(1 << 24)
        out[k] = (torch.sin(r / 10.0) + u / 100.0).to(torch.float32)
    return out


def quality(v, w):
    """error = rms(v-w)/rms(v), SNR = -20 log10(error): Test_With_Generated_Input.cpp:98-114."""
    v = np.asarray(v, dtype=np.float64).ravel()
    w = np.asarray(w, dtype=np.float64).ravel()
    num = np.sqrt(np.mean((v - w) ** 2))
    den = np.sqrt(np.mean(v ** 2))
    err = num / den if den > 0 else num
    snr = -20.0 * np.log10(err) if err > 0 else float("inf")
    return err, snr
