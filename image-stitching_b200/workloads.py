"""Benchmark / parity workloads of BASELINE.json `configs`.

Homographies are forward (source pixel -> canvas pixel), row-major, exactly as they appear
in stitch.conf.xml (/root/reference/tools/calibrate/main.cpp:497-522 writes them,
/root/reference/tools/stitch/main.cpp:60-73 reads them).  Values are the authoritative
ones of SURVEY.md Appendix B.2 / B.1, verbatim.
"""
import numpy as np

_R23_EVEN = [0.009495859491008554, 1.0382736317291528, -59.197572490221546,
             2.0690585458024008e-06, -1.574483005320176e-06, 0.9977278492378052]
_R23_ODD = [-0.009519391604875454, 1.0416443039862457, -26.32861642378475,
            -2.090840077932041e-06, 1.5454409624806558e-06, 1.0023453367101505]
_ROW1_4K = [
    [1.0437883446452338, -0.010162597785286163, -149.89801611073972],
    [1.0321064065523367, 0.013076314139887218, 1767.4332801484354],
    [1.0517335294611148, -0.01620861252571564, 3681.3769249624324],
    [1.0240775806530777, 0.019010807435812935, 5616.439373115412],
    [1.0596787142769961, -0.022254627266145115, 7512.651866035605],
    [1.0160487547538186, 0.024945300731738656, 9465.44546608239],
    [1.0676238990928772, -0.028300642006574592, 11343.926807108777],
    [1.0080199288545595, 0.030879794027664376, 13314.451559049368],
]

_H_1080 = [
    [1.041881500289422, -0.008711554247583088, -73.90200098415016,
     0.008378567876275259, 1.0391238525520257, -28.80030553931806,
     2.0690585458024008e-06, -1.574483005320176e-06, 0.9988639246189026],
    [1.0340333247681588, 0.011652035748865048, 1842.6359089181803,
     -0.008390337962792151, 1.0408097658665063, -13.612549123632903,
     -2.090840077932041e-06, 1.5454409624806558e-06, 1.0011726683550752],
]

# calibrate --mode=extend -b=3x4 on test/Inputs/1a.jpg, 1b.jpg
_H_CFG1 = [
    [1.1120443247359568, 0.2321184108605204, 192.93634269880312,
     -0.7385466102701779, 2.359969315805939, 2227.1748206665548,
     -0.00013831019805894855, 0.00018804770166660635, 1.0],
    [1.1355734419545565, 0.16166738072381795, 0.000368944293145294,
     -0.5899936756162424, 2.554407720031592, 3521.8550208170154,
     -9.468468147649584e-05, 0.00018031156578367973, 1.0],
]


def _h4k(i):
    return _ROW1_4K[i] + (_R23_EVEN if i % 2 == 0 else _R23_ODD)


class Workload:
    def __init__(self, name, src_wh, canvas_wh, H, algorithmic_bytes=None):
        self.name = name
        self.n_cams = len(H)
        self.src_wh = src_wh              # (w, h) of every camera frame
        self.canvas_wh = canvas_wh        # (W, H) of the panorama
        self.H = np.asarray(H, dtype=np.float64).reshape(self.n_cams, 3, 3)
        # W*H*3 + 3*sum_i U_i  (4.13 border rule)
        self.algorithmic_bytes = algorithmic_bytes

    @property
    def src_bytes(self):
        return self.n_cams * self.src_wh[0] * self.src_wh[1] * 3

    @property
    def canvas_bytes(self):
        return self.canvas_wh[0] * self.canvas_wh[1] * 3


WORKLOADS = {
    # configs[0]: two chessboard photos (3016x4032) -> 6085x8353
    # (56 % of the canvas is covered; the uncovered part is stored as zeros, so it counts as canvas bytes)
    "cfg1": Workload("cfg1", (3016, 4032), (6085, 8353), _H_CFG1, 196608033),
    # configs[1]: 2 x 1080p -> 3840x1080
    "cfg2": Workload("cfg2", (1920, 1080), (3840, 1080), _H_1080, 23962032),
    # configs[2], [3]: 4 x 4K -> 7680x2160 (headline)
    "cfg3": Workload("cfg3", (3840, 2160), (7680, 2160), [_h4k(i) for i in range(4)], 95836773),
    # configs[4]: 8 x 4K ring -> 15360x2160
    "cfg5": Workload("cfg5", (3840, 2160), (15360, 2160), [_h4k(i) for i in range(8)], 191662263),
}


def get(name):
    return WORKLOADS[name]


# ---- synthetic frame content ---------------------------------------
def gen_frame_numpy(w, h, cam, frame, kind="noise"):
    """Camera `cam`, frame `frame`, packed BGR u8.  kind: 'noise' (hash) or 'smooth' (ramp)."""
    x = np.arange(w, dtype=np.uint32)[None, :, None]
    y = np.arange(h, dtype=np.uint32)[:, None, None]
    c = np.arange(3, dtype=np.uint32)[None, None, :]
    if kind == "smooth":
        return ((x + 2 * y + 64 * c + np.uint32(7 * frame + 31 * cam)) & 255).astype(np.uint8)
    with np.errstate(over="ignore"):
        v = (x * np.uint32(0x9E3779B1)) ^ (y * np.uint32(0x85EBCA77)) ^ (c * np.uint32(0xC2B2AE3D))
        v = v ^ (np.uint32(cam * 131 + frame) * np.uint32(0x27D4EB2F))
        v ^= v >> 15
        v *= np.uint32(0x2C1B3C6D)
        v ^= v >> 12
        v *= np.uint32(0x297A2D39)
        v ^= v >> 15
    return (v >> 24).astype(np.uint8)


def gen_frame_torch(w, h, cam, frame, kind="noise", device="cuda"):
    """Same content generated on the device (uint32 arithmetic emulated in int64)."""
    import torch
    m = 0xFFFFFFFF
    x = torch.arange(w, dtype=torch.int64, device=device)[None, :, None]
    y = torch.arange(h, dtype=torch.int64, device=device)[:, None, None]
    c = torch.arange(3, dtype=torch.int64, device=device)[None, None, :]
    if kind == "smooth":
        return ((x + 2 * y + 64 * c + (7 * frame + 31 * cam)) & 255).to(torch.uint8).contiguous()
    v = ((x * 0x9E3779B1) & m) ^ ((y * 0x85EBCA77) & m) ^ ((c * 0xC2B2AE3D) & m)
    v = v ^ (((cam * 131 + frame) * 0x27D4EB2F) & m)
    v = v ^ (v >> 15)
    v = (v * 0x2C1B3C6D) & m
    v = v ^ (v >> 12)
    v = (v * 0x297A2D39) & m
    v = v ^ (v >> 15)
    return (v >> 24).to(torch.uint8).contiguous()
