"""Case tables shared by tests/golden/make_golden.py (which runs cv2) and the tests (which
must not need cv2 or /root/reference)."""
import numpy as np

MODE_SGBM, MODE_HH, MODE_3WAY, MODE_HH4 = 0, 1, 2, 3


def _p(minD, D, bs, mode, uniq=10, spw=100, spr=32, d12=1, cap=63, p1=None, p2=None):
    return dict(minDisparity=minD, numDisparities=D, blockSize=bs,
                P1=8 * bs * bs if p1 is None else p1, P2=32 * bs * bs if p2 is None else p2,
                disp12MaxDiff=d12, preFilterCap=cap, uniquenessRatio=uniq,
                speckleWindowSize=spw, speckleRange=spr, mode=mode)


SGBM_CASES = [
    dict(H=60, W=96, seed=11, prm=_p(0, 16, 3, MODE_SGBM)),
    dict(H=72, W=130, seed=12, prm=_p(0, 32, 5, MODE_SGBM, spw=50, spr=2)),
    dict(H=64, W=141, seed=13, prm=_p(0, 64, 5, MODE_HH)),
    dict(H=81, W=131, seed=14, prm=_p(0, 48, 7, MODE_3WAY)),
    dict(H=66, W=110, seed=15, prm=_p(-8, 32, 3, MODE_SGBM, uniq=5, spw=0)),
    dict(H=70, W=120, seed=16, prm=_p(4, 32, 5, MODE_HH, uniq=15, d12=2, cap=31)),
    dict(H=57, W=200, seed=17, prm=_p(0, 128, 5, MODE_SGBM)),
    dict(H=48, W=120, seed=18, prm=_p(0, 80, 7, MODE_3WAY, uniq=0, spw=0)),     # D not a power of two
    dict(H=64, W=100, seed=19, prm=_p(0, 32, 9, MODE_SGBM, p1=1944, p2=7776)),   # S saturates at 32767
    dict(H=40, W=400, seed=20, prm=_p(0, 320, 7, MODE_3WAY)),                    # the reference's own D / bs
    # MODE_HH4 (4 axis-aligned paths; the rows within blockSize/2 of the bottom edge cost 0)
    dict(H=64, W=141, seed=21, prm=_p(0, 64, 5, MODE_HH4)),
    dict(H=70, W=120, seed=22, prm=_p(-8, 32, 9, MODE_HH4, uniq=5, d12=2, cap=31, spw=0)),
    dict(H=33, W=230, seed=23, prm=_p(4, 144, 3, MODE_HH4, uniq=15)),
    # 3WAY without a uniqueness test: cv2's 8-lane argmin breaks cost ties its own way (later d within a lane, then
    # the lowest lane result); both pairs contain such ties (d = 28 / 60 and d = 5 / 21).  Also minDisparity > 0.
    dict(H=43, W=151, seed=519, prm=_p(7, 64, 7, MODE_3WAY, uniq=0, spw=0)),
    dict(H=41, W=159, seed=520, prm=_p(16, 48, 5, MODE_3WAY, uniq=0, spw=0)),
    dict(H=67, W=140, seed=91, prm=_p(4, 32, 5, MODE_3WAY)),
]

# 3-channel input (cv2 sums the BT cost over the colour channels); P1/P2 as the reference's
# 8*3*bs^2 / 32*3*bs^2 (depth_map.py:898-899)
SGBM_COLOR_CASES = [
    dict(H=60, W=120, seed=24, prm=_p(0, 32, 5, MODE_SGBM, p1=600, p2=2400)),
    dict(H=52, W=150, seed=25, prm=_p(0, 64, 3, MODE_HH, p1=216, p2=864, uniq=5)),
    dict(H=64, W=140, seed=26, prm=_p(0, 48, 7, MODE_3WAY, p1=1176, p2=4704)),
    dict(H=40, W=100, seed=27, prm=_p(-8, 16, 5, MODE_HH4, p1=600, p2=2400, cap=31)),
]


def color_pair(c):
    """Interleaved BGR pair of a SGBM_COLOR_CASES entry: per-channel gains plus independent seeded noise."""
    from stereovision_b200.synth import make_pair, to_bgr
    l, r, _ = make_pair(c["H"], c["W"], max(c["prm"]["numDisparities"], 16), seed=c["seed"])
    rng = np.random.default_rng(5000 + c["seed"])
    out = []
    for g in (l, r):
        bgr = to_bgr(g).astype(np.int16) + rng.integers(-12, 13, g.shape + (3,))
        out.append(np.clip(bgr, 0, 255).astype(np.uint8))
    return out[0], out[1]


BM_CASES = [
    dict(H=60, W=120, seed=31, numDisparities=16, blockSize=9),
    dict(H=72, W=150, seed=32, numDisparities=64, blockSize=15),
    dict(H=65, W=131, seed=33, numDisparities=32, blockSize=21),
    dict(H=64, W=128, seed=34, numDisparities=32, blockSize=7, set={"MinDisparity": -4, "UniquenessRatio": 0}),
    dict(H=64, W=140, seed=35, numDisparities=48, blockSize=11,
         set={"TextureThreshold": 50, "SpeckleWindowSize": 30, "SpeckleRange": 4}),
    # disp12MaxDiff >= 0 (cv2.validateDisparity inside StereoBM)
    dict(H=81, W=98, seed=36, numDisparities=32, blockSize=9, noise=25,
         set={"Disp12MaxDiff": 1, "UniquenessRatio": 0, "TextureThreshold": 0}),
    dict(H=70, W=160, seed=37, numDisparities=64, blockSize=15, noise=20,
         set={"Disp12MaxDiff": 0, "MinDisparity": -16, "SpeckleWindowSize": 30, "SpeckleRange": 4}),
    dict(H=64, W=141, seed=38, numDisparities=48, blockSize=7, noise=12, set={"Disp12MaxDiff": 3}),
    # PREFILTER_NORMALIZED_RESPONSE (type 0), default window 9 and others
    dict(H=70, W=150, seed=39, numDisparities=32, blockSize=9, set={"PreFilterType": 0}),
    dict(H=61, W=133, seed=40, numDisparities=48, blockSize=15,
         set={"PreFilterType": 0, "PreFilterSize": 5, "PreFilterCap": 63, "MinDisparity": -4}),
    dict(H=66, W=120, seed=41, numDisparities=16, blockSize=7, noise=20,
         set={"PreFilterType": 0, "PreFilterSize": 31, "PreFilterCap": 10, "Disp12MaxDiff": 1}),
    # minDisparity > 0: cv2 evaluates minDisparity positions past the row end; the spill of the last ROI row lands
    # in the first columns of row H - blockSize/2 and survives (oracle/bm_oracle.c)
    dict(H=75, W=149, seed=42, numDisparities=16, blockSize=5,
         set={"MinDisparity": 14, "UniquenessRatio": 0, "TextureThreshold": 0}),
    dict(H=60, W=196, seed=43, numDisparities=48, blockSize=9, noise=12,
         set={"MinDisparity": 11, "UniquenessRatio": 0, "SpeckleWindowSize": 6, "SpeckleRange": 8}),
    dict(H=64, W=150, seed=44, numDisparities=32, blockSize=7, noise=20,
         set={"MinDisparity": 5, "Disp12MaxDiff": 1, "SpeckleWindowSize": 20, "SpeckleRange": 2}),
]


def synth_calibration(W, H):
    """Synthetic stereo calibration turned into rectification maps with the
    reference's own recipe (depth_map.py:613-638).  Needs cv2: only the golden generator and
    cv2-gated tests call this."""
    import cv2
    fx = 0.78 * W
    K = np.array([[fx, 0, W / 2 + 3], [0, fx, H / 2 - 2], [0, 0, 1]], np.float64)
    dl = np.array([-0.10, 0.02, 1e-3, -1e-3, 0], np.float64)
    dr = np.array([-0.09, 0.015, -1e-3, 1e-3, 0], np.float64)
    R, _ = cv2.Rodrigues(np.array([0.01, -0.02, 0.005]))
    T = np.array([[-6.0], [0.05], [0.02]])
    R1, R2, P1, P2, Q, _, _ = cv2.stereoRectify(K, dl, K, dr, (W, H), R, T, alpha=0)
    m1l, m2l = cv2.initUndistortRectifyMap(K, dl, R1, P1, (W, H), cv2.CV_16SC2)
    m1r, m2r = cv2.initUndistortRectifyMap(K, dr, R2, P2, (W, H), cv2.CV_16SC2)
    return dict(mtx_left=K, dist_left=dl, mtx_right=K, dist_right=dr, R=R, T=T, img_size=(W, H),
                left_map1=m1l, left_map2=m2l, right_map1=m1r, right_map2=m2r, Q=Q)


# Fusion chain (fused_depth_map.py): stereo (min_disp, num_disp, window) as the reference's scaled
# settings give them, plus what the golden generator perturbs.
FUSION_CASES = [
    dict(H=96, W=160, seed=41, min_disp=0, num_disp=48, window=5, moving=True, params={}),
    dict(H=120, W=176, seed=42, min_disp=0, num_disp=64, window=7, moving=False, params={}),
    dict(H=90, W=150, seed=43, min_disp=0, num_disp=32, window=5, moving=True,
         params=dict(stereo_weight=1.1, midas_fill_weight=0.6, flow_fill_weight=0.3,
                     stereo_conf_threshold=0.4, flow_hole_threshold=25)),
    dict(H=96, W=144, seed=44, min_disp=0, num_disp=48, window=5, moving=True, params={}, occluded="right"),
]


def fusion_case_inputs(c):
    """(left BGR, right BGR, MiDaS-like f32, flow-depth f32) of a FUSION_CASES entry, from its seed."""
    from stereovision_b200.synth import make_pair, make_aux_depths, to_bgr
    l, r, gt = make_pair(c["H"], c["W"], max(c["num_disp"], 16), seed=c["seed"])
    if c.get("occluded") == "right":
        r = np.full_like(r, 22) + (r // 64)
    midas, flow = make_aux_depths(gt, c["seed"])
    return to_bgr(l), to_bgr(r), midas, flow


# depth_map.py drop-in (apply_stereo_rectification + create_depth_map with the script's own
# MIN_DISP=0 / NUM_DISP=320 / WINDOW_SIZE=7 / 3WAY): W must exceed 320.
DROPIN_CASE = dict(H=150, W=520, seed=51)


# the code around the mono estimators (fused_depth_map.py:1134-1158, 1386-1403).
# pred: shape of the network output that is upsampled to the frame; flat: constant prediction / frame.
MONO_CASES = [
    dict(H=120, W=176, seed=61, pred=(48, 64)),
    dict(H=97, W=131, seed=62, pred=(24, 32)),
    dict(H=64, W=80, seed=63, pred=(96, 128), flat=True),          # the degenerate ranges
    dict(H=104, W=136, seed=64, pred=(150, 200)),                    # prediction larger than the frame
]


def mono_case_inputs(c):
    """(frame BGR u8, network prediction f32 [h,w]) of a MONO_CASES entry, from its seed."""
    from stereovision_b200.synth import make_pair, to_bgr
    l, _, _ = make_pair(c["H"], c["W"], 16, seed=c["seed"])
    rng = np.random.default_rng(2000 + c["seed"])
    h, w = c["pred"]
    if c.get("flat"):
        return to_bgr(np.full_like(l, 77)), np.full((h, w), 0.5, np.float32)
    yy, xx = np.mgrid[0:h, 0:w]
    pred = 400.0 + 300.0 * np.sin(xx / 9.0 + rng.uniform(0, 3)) * np.cos(yy / 7.0) + rng.normal(0, 25.0, (h, w))
    return to_bgr(l), pred.astype(np.float32)


def mono_flow_frames(c, n=4):
    """n BGR frames of a camera panning over the textured image of a MONO_CASES entry (integer shifts plus
    a little noise), for OpticalFlowDepthEstimator."""
    from stereovision_b200.synth import make_pair, to_bgr
    l, _, _ = make_pair(c["H"] + 16, c["W"] + 32, 16, seed=c["seed"])
    rng = np.random.default_rng(3000 + c["seed"])
    frames = []
    for k in range(n):
        crop = l[2 * k:2 * k + c["H"], 3 * k:3 * k + c["W"]].astype(np.int16)
        crop[c["H"] // 3:c["H"] // 3 + 20, c["W"] // 2:c["W"] // 2 + 30] = l[8:28, 40 + 6 * k:70 + 6 * k]   # a nearer object
        frames.append(to_bgr(np.clip(crop + rng.integers(-2, 3, crop.shape), 0, 255).astype(np.uint8)))
    return frames


# SURVEY 8(f) N4 (display-side reductions): W must exceed depth_map.py's NUM_DISP = 320.
DISPLAY_CASE = dict(H=120, W=480, seed=71, fx=374.4, min_depth=0.2, max_depth=40.0)


def display_case_inputs():
    """(MiDaS-like f32 map stretched beyond 0..255, u8 map for the legend) of DISPLAY_CASE."""
    from stereovision_b200.synth import make_pair, make_aux_depths
    c = DISPLAY_CASE
    _, _, gt = make_pair(c["H"], c["W"], 64, seed=c["seed"])
    midas, flow = make_aux_depths(gt, c["seed"])
    wide = (midas.astype(np.float32) - np.float32(midas.mean())) * np.float32(1.7) + np.float32(120.0)
    return wide, np.where(flow > 0, np.clip(flow, 90, 255), 0).astype(np.uint8)


def dense_disparity(c=None):
    """A float32 disparity field covering the multiples of 1/16 in [-1, 320) (what compute()/16 can hold)."""
    c = c or DISPLAY_CASE
    rng = np.random.default_rng(4000 + c["seed"])
    d = rng.integers(-16, 320 * 16, (c["H"], c["W"])).astype(np.float32) / np.float32(16.0)
    d[:, :40] = rng.integers(1, 16 * 16, (c["H"], 40)).astype(np.float32) / np.float32(16.0)     # plenty of near range
    return d


# cv2.resize(INTER_LINEAR) (fused_depth_map.py:473-476, depth_map.py:809-811, fused_depth_map.py:1214-1217):
# (source H, W, channels or "f32", destination (W, H)).  The first is the reference's own case: a 640x480 BGR frame
# to int(640 * 0.33) x int(480 * 0.33) (fused_depth_map.py:385-386).
RESIZE_CASES = [
    dict(H=480, W=640, ch=3, dsize=(211, 158), seed=81),
    dict(H=97, W=131, ch=1, dsize=(200, 150), seed=82),       # upscale, odd sizes: the first / last rows split weights
    dict(H=120, W=160, ch=3, dsize=(80, 60), seed=83),        # exact 2x down
    dict(H=77, W=101, ch=3, dsize=(33, 41), seed=84),
    dict(H=64, W=48, ch=1, dsize=(48, 64), seed=85),          # same size: a copy
    dict(H=96, W=128, ch="f32", dsize=(211, 158), seed=86),   # mono map (384-ish network output) to the stereo shape
    dict(H=150, W=200, ch="f32", dsize=(131, 97), seed=87),
]


def resize_case_input(c):
    rng = np.random.default_rng(c["seed"])
    if c["ch"] == "f32":
        yy, xx = np.mgrid[0:c["H"], 0:c["W"]]
        return (400.0 + 300.0 * np.sin(xx / 9.0) * np.cos(yy / 7.0) + rng.normal(0, 25.0, (c["H"], c["W"]))).astype(np.float32)
    shape = (c["H"], c["W"]) if c["ch"] == 1 else (c["H"], c["W"], 3)
    return rng.integers(0, 256, shape, dtype=np.uint8)


# Calibration-file path (depth_map.py:591-668, fused_depth_map.py:307-442): synthetic .pkl at this size, the
# scaled loader at the reference's default PROCESSING_SCALE = 0.33 and two more.
CALIB_CASE = dict(W=320, H=240, scales=(1.0, 0.5, 0.33))


# BASELINE.json config 5 at its own size: 1080p stereo depth (D = 256) + two synthetic mono maps, normal and with
# the right camera covered (tests/golden/make_golden_fusion_1080p.py).
FUSION_1080P = dict(H=1080, W=1920, seed=45, min_disp=0, num_disp=256, window=5, moving=True, params={})
