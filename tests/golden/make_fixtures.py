"""Generate the committed fixtures under tests/golden/ from the reference tree.

Run HERE (the authoring container, where /root/reference exists):
    python tests/golden/make_fixtures.py
Nothing in tests/, smoke() or bench.py reads /root/reference at run time; they read the
.npz files this script writes.

  raytrace_b200/assets/{pawn.obj,bunny.obj,earthmap.jpg} — the reference's input files, copied as package data (what the
                                       product reads: .obj text through rt_obj_parse, the JPEG through readImage)
  assets/pawn.npz, assets/bunny.npz  — images/*.obj parsed by the pure-Python restatement of
                                       parseObj (oracle/obj_oracle.py; Geometry.hs:207-285)
  assets/earthmap.npz                — images/earthmap.jpg decoded by PIL to 8-bit sRGB
                                       (readImage's decode step, Ray.hs:241-245; the sRGB→linear
                                       step is applied at load time)
  png/<name>.npz                     — the reference's golden renders, decoded to 8-bit RGB:
                                       example_image (README scene, seed 100, 50 spp, writeImage),
                                       cornell_box_noisy / cornell_box_redirect (200 spp, writeImageSqrt),
                                       pawn_demo (500x500, 400 spp, depth 20, writeImage),
                                       demo2 (800x800, scene from mkStdGen 1234, writeImageSqrt; spp / depth unrecorded),
                                       demo1 (1200x675, small spheres from newStdGen - only the camera, the ground, the sky and the
                                       three big spheres are deterministic; writeImageSqrt)
"""
import os
import sys

import numpy as np
from PIL import Image

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
from oracle.obj_oracle import parse_obj  # noqa: E402

REF = "/root/reference"


def main():
    import shutil
    pkg = os.path.join(ROOT, "raytrace_b200", "assets")
    os.makedirs(pkg, exist_ok=True)
    for f in ("pawn.obj", "bunny.obj", "earthmap.jpg"):
        shutil.copyfile(f"{REF}/images/{f}", os.path.join(pkg, f))
    os.makedirs(os.path.join(HERE, "assets"), exist_ok=True)
    os.makedirs(os.path.join(HERE, "png"), exist_ok=True)
    for name in ("pawn", "bunny"):
        with open(f"{REF}/images/{name}.obj") as f:
            v, t, fc = parse_obj(f.read())
        np.savez_compressed(os.path.join(HERE, "assets", f"{name}.npz"), verts=v, uvs=t, faces=fc)
        print(name, v.shape, t.shape, fc.shape)
    img = np.asarray(Image.open(f"{REF}/images/earthmap.jpg").convert("RGB"), dtype=np.uint8)
    np.savez_compressed(os.path.join(HERE, "assets", "earthmap.npz"), rgb8=img)
    print("earthmap", img.shape)
    for name in ("example_image", "cornell_box_noisy", "cornell_box_redirect", "pawn_demo", "demo2", "demo1"):
        a = np.asarray(Image.open(f"{REF}/{name}.png").convert("RGB"), dtype=np.uint8)
        np.savez_compressed(os.path.join(HERE, "png", f"{name}.npz"), rgb8=a)
        print(name, a.shape)


if __name__ == "__main__":
    main()
