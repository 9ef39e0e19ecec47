#!/usr/bin/env python
"""What the e2e path's copies cost by themselves, pattern by pattern, at cfg3's geometry: 4 frames of 3840x2160x3
of which 47 % of every row is uploaded (2-D copies), one 7680x2160x3 panorama downloaded.  Explains the distance
between `e2e.value` and `e2e.host_ceiling_fps` (which uses 1-D copies of the same byte counts).
usage: python tools/copy_patterns.py"""
import json, time
from cuda import cudart

def ck(r):
    if isinstance(r, tuple):
        if r[0] != cudart.cudaError_t.cudaSuccess: raise RuntimeError(str(r[0]))
        return r[1] if len(r) == 2 else r[1:]
    if r != cudart.cudaError_t.cudaSuccess: raise RuntimeError(str(r))

ck(cudart.cudaSetDevice(0))
SW, SH, W, H, N = 3840, 2160, 7680, 2160, 4
row = SW * 3
part = int(0.47 * row) // 64 * 64
DEPTH = 3
h_frames = [[ck(cudart.cudaHostAlloc(row * SH, cudart.cudaHostAllocDefault)) for _ in range(N)] for _ in range(DEPTH)]
d_frames = [[ck(cudart.cudaMalloc(row * SH)) for _ in range(N)] for _ in range(DEPTH)]
h_out = [ck(cudart.cudaHostAlloc(W * 3 * H, cudart.cudaHostAllocDefault)) for _ in range(DEPTH)]
d_out = [ck(cudart.cudaMalloc(W * 3 * H)) for _ in range(DEPTH)]
s_up = ck(cudart.cudaStreamCreateWithFlags(cudart.cudaStreamNonBlocking))
s_dn = ck(cudart.cudaStreamCreateWithFlags(cudart.cudaStreamNonBlocking))
H2D, D2H = cudart.cudaMemcpyKind.cudaMemcpyHostToDevice, cudart.cudaMemcpyKind.cudaMemcpyDeviceToHost

def up_1d(k):
    for i in range(N): ck(cudart.cudaMemcpyAsync(d_frames[k][i], h_frames[k][i], part * SH, H2D, s_up))
def up_2d(k):
    for i in range(N): ck(cudart.cudaMemcpy2DAsync(d_frames[k][i], row, h_frames[k][i], row, part, SH, H2D, s_up))
def up_full(k):
    for i in range(N): ck(cudart.cudaMemcpyAsync(d_frames[k][i], h_frames[k][i], row * SH, H2D, s_up))
def dn_1d(k): ck(cudart.cudaMemcpyAsync(h_out[k], d_out[k], W * 3 * H, D2H, s_dn))
def dn_2d(k): ck(cudart.cudaMemcpy2DAsync(h_out[k], W * 3, d_out[k], W * 3, W * 3, H, D2H, s_dn))

def run(up, dn, frames=60):
    for f in range(frames + 6):
        if f == 6:
            ck(cudart.cudaStreamSynchronize(s_up)); ck(cudart.cudaStreamSynchronize(s_dn)); t0 = time.perf_counter()
        if up: up(f % DEPTH)
        if dn: dn(f % DEPTH)
    ck(cudart.cudaStreamSynchronize(s_up)); ck(cudart.cudaStreamSynchronize(s_dn))
    return frames / (time.perf_counter() - t0)

res = {"h2d_bytes_partial": part * SH * N, "d2h_bytes": W * 3 * H}
for name, up, dn in (("up1d+dn1d (what host_ceiling measures)", up_1d, dn_1d), ("up2d+dn2d (what bs_submit issues)", up_2d, dn_2d),
                     ("up2d+dn1d", up_2d, dn_1d), ("up2d only", up_2d, None), ("up1d only", up_1d, None), ("dn1d only", None, dn_1d),
                     ("dn2d only", None, dn_2d), ("full frames 1d + dn1d", up_full, dn_1d)):
    res[name] = round(run(up, dn), 1)
print(json.dumps(res, indent=1))
