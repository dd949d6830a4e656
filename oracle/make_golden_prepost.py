"""Generates tests/golden/prepost.npz from the UNMODIFIED reference driver helpers.

The reference drivers are scripts with heavy top-level imports (imageio, a CUDA extension, argparse mains), so the helper
functions are pulled out of their source files with `ast` and executed here exactly as written:
    python oracle/make_golden_prepost.py        (needs /root/reference; run in the build container only)
"""
import ast
import os
import struct
import sys
import tempfile

import cv2
import numpy as np
import torch
import torch.nn.functional as F

REF = "/root/reference"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def extract(path, names):
    src = open(os.path.join(REF, path)).read()
    tree = ast.parse(src)
    ns = {"np": np, "torch": torch, "F": F, "cv2": cv2, "struct": struct, "sys": sys}
    for node in tree.body:
        if isinstance(node, ast.FunctionDef) and node.name in names:
            exec(compile(ast.Module([node], []), path, "exec"), ns)
    missing = [n for n in names if n not in ns]
    assert not missing, (path, missing)
    return ns


def main():
    vid = extract("pwc_extract_flow_video.py", ["frame_to_tensor", "pad_to_multiple_of_64", "unpad"])
    kit = extract("inference_kitti.py", ["pad_to_64", "unpad", "flow_resize", "epe_metric", "fl_all_metric", "load_flow_kitti_png"])
    scr = extract("script_pwc.py", ["writeFlowFile"])
    ext = extract("pwc_extract_flow.py", ["write_flo"])
    dpr = extract("data_processing.py", ["read_flow"])
    g = np.random.default_rng(7)
    out = {}
    # ---- f1 pre: two BGR uint8 frames 37x50 -> [1,6,64,64]
    fr1 = g.integers(0, 256, (37, 50, 3), dtype=np.uint8)
    fr2 = g.integers(0, 256, (37, 50, 3), dtype=np.uint8)
    t1, t2 = vid["frame_to_tensor"](fr1)[None], vid["frame_to_tensor"](fr2)[None]
    t1p, t2p, ph, pw = vid["pad_to_multiple_of_64"](t1, t2)
    out.update(pre_fr1=fr1, pre_fr2=fr2, pre_x=torch.cat([t1p, t2p], 1).numpy(), pre_pad=np.array([ph, pw]))
    # all 256 byte values through astype(float32)/255.0
    ramp = np.arange(256, dtype=np.uint8).reshape(1, 256, 1).repeat(3, 2)
    out["pre_ramp"] = vid["frame_to_tensor"](ramp).numpy()
    # ---- f1 post: quarter-resolution flow [2,2,16,16] of a 37x50 frame pair: unpad (reference quirk: image-pixel pads cut
    #      from the quarter-res flow) then flow_resize to 37x50
    fl = torch.from_numpy(g.standard_normal((2, 2, 16, 16)).astype(np.float32) * 3)
    un = kit["unpad"](fl, 5, 3)
    out.update(post_flow=fl.numpy(), post_unpad_shape=np.array(un.shape), post_resized=kit["flow_resize"](un.clone(), 37, 50).numpy(),
               post_resized_full=kit["flow_resize"](fl.clone(), 64, 64).numpy(),
               post_video_unpad_shape=np.array(vid["unpad"](fl, 5, 3).shape))
    # ---- f3 metrics
    pred = (g.standard_normal((23, 31, 2)) * 4).astype(np.float32)
    gt = (g.standard_normal((23, 31, 2)) * 4 + 1).astype(np.float32)
    gt[:4] *= 30                                             # large ground-truth vectors: the 5% branch of the threshold
    valid = g.random((23, 31)) > 0.3
    out.update(met_pred=pred, met_gt=gt, met_valid=valid,
               met_epe=np.array([kit["epe_metric"](pred, gt, valid), kit["epe_metric"](pred, gt, None)]),
               met_fl=np.array([kit["fl_all_metric"](pred, gt, valid), kit["fl_all_metric"](pred, gt, None)]))
    # ---- f2 formats
    with tempfile.TemporaryDirectory() as d:
        uv = (g.standard_normal((5, 7, 2)) * 10).astype(np.float32)
        scr["writeFlowFile"](os.path.join(d, "a.flo"), uv)
        ext["write_flo"](os.path.join(d, "b.flo"), uv)
        a, b = open(os.path.join(d, "a.flo"), "rb").read(), open(os.path.join(d, "b.flo"), "rb").read()
        assert a == b
        out.update(flo_uv=uv, flo_bytes=np.frombuffer(a, np.uint8), flo_read=dpr["read_flow"](os.path.join(d, "a.flo")))
        # KITTI 16-bit PNG as the reference READS it: a file written by cv2 (B,G,R order on disk = valid, v, u)
        rgb = g.integers(0, 65536, (6, 9, 3), dtype=np.uint16)
        rgb[..., 2] = g.integers(0, 2, (6, 9))
        p = os.path.join(d, "k.png")
        cv2.imwrite(p, rgb[..., ::-1].copy())
        fl_k, va_k = kit["load_flow_kitti_png"](p)
        out.update(kitti_rgb=rgb, kitti_png=np.frombuffer(open(p, "rb").read(), np.uint8), kitti_flow=fl_k, kitti_valid=va_k)
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "prepost.npz"), **out)
    print("wrote tests/golden/prepost.npz:", {k: getattr(v, "shape", None) for k, v in out.items()})


if __name__ == "__main__":
    main()
