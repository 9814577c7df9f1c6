"""tests/golden/make_assets.py — writes this repository's OWN small asset set (not derived from the reference):
a two-mesh OBJ with UVs (a wedge and a quad), its MTL, two 32x32 RGBA PNGs and scene files in the reference's
grammar.  Used by the host-class tests (parser, OBJ/MTL/PNG import) on any machine."""
import os
import struct
import zlib

import numpy as np

HERE = os.path.join(os.path.dirname(os.path.abspath(__file__)), "assets")


def write_png(path, rgba):
    h, w, _ = rgba.shape
    raw = b"".join(b"\x00" + rgba[y].tobytes() for y in range(h))

    def chunk(t, d):
        return struct.pack(">I", len(d)) + t + d + struct.pack(">I", zlib.crc32(t + d) & 0xFFFFFFFF)
    with open(path, "wb") as f:
        f.write(b"\x89PNG\r\n\x1a\n" + chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, 8, 6, 0, 0, 0)) +
                chunk(b"IDAT", zlib.compress(raw, 9)) + chunk(b"IEND", b""))


def main():
    os.makedirs(HERE, exist_ok=True)
    y, x = np.mgrid[0:32, 0:32]
    a = np.stack([(x * 8) % 256, (y * 8) % 256, ((x ^ y) * 8) % 256, 255 - x * 4], -1).astype(np.uint8)
    b = np.stack([255 * ((x // 4 + y // 4) % 2), 64 + x * 4, 255 - y * 6, np.full_like(x, 200)], -1).astype(np.uint8)
    write_png(os.path.join(HERE, "tex_a.png"), a)
    write_png(os.path.join(HERE, "tex_b.png"), b)
    with open(os.path.join(HERE, "shapes.mtl"), "w") as f:
        f.write("newmtl wedge_mat\nKd 0.8 0.8 0.8\nmap_Kd /somewhere/else/on/another/machine/tex_a.png\n\n"
                "newmtl quad_mat\nKd 0.5 0.5 0.5\nmap_Kd tex_b.png\n")
    with open(os.path.join(HERE, "shapes.obj"), "w") as f:
        f.write("""# two meshes: a closed wedge (CCW, outward) and a single quad
mtllib shapes.mtl
o Wedge
v -1.0 -1.0 -1.0
v 1.0 -1.0 -1.0
v 1.0 -1.0 1.0
v -1.0 -1.0 1.0
v -1.0 1.0 -1.0
v 1.0 1.0 -1.0
vt 0.0 0.0
vt 1.0 0.0
vt 1.0 1.0
vt 0.0 1.0
vt 0.5 0.25
usemtl wedge_mat
f 1/1 2/2 3/3 4/4
f 1/1 5/4 6/3 2/2
f 4/1 3/2 6/3 5/4
f 1/1 4/2 5/5
f 2/1 6/5 3/2
o Quad
v -2.0 0.5 3.0
v 2.0 0.5 3.0
v 2.0 2.5 3.0
v -2.0 2.5 3.0
usemtl quad_mat
f -4/1 -1/4 -2/3 -3/2
""")
    with open(os.path.join(HERE, "shapes.scene"), "w") as f:
        f.write("""# repository-owned test scene in the reference's grammar
MATERIALS:
textured, (1, 1, 1), 0.9        #0
diffuse, (0.7, 0.7, 0.75), 1    #1
light, (1, 1, 1), 0             #2
dielectric, (1, 1, 1), 1.4      #3
reflective, (1, 0.9, 0.8), 0.8  #4
refractive, (1, 1, 1), 1.2      #5

SPHERES:
(0, -150, 5), 80, 2
(-2.5, 0.3, 6), 0.9, 3
(2.6, .4, 7), +1.1, 4
(0.2, 0.8, 3.2), 0.45, 5

PLANES:
(0, 1.5, 0), (0, 1, 0), 1

LENSES:
(-1, -0.5, 4), (0, 0, 1), 3, 3, 0.8, 5

MODELS:
translate: (0.5, 0.4, 8)
rotate: 40, (0, 1, 0)
scale: (0.8, 0.8, 0.8)
load: "shapes.obj", 0
""")
    bad = {
        "bad_material.scene": "MATERIALS:\nshiny, (1, 1, 1), 1\n",
        "bad_vector.scene": "MATERIALS:\ndiffuse, (1, 1, 1), 1\nSPHERES:\n(0 , 0, 3), 1, 0\n",
        "bad_uint.scene": "MATERIALS:\ndiffuse, (1, 1, 1), 1\nSPHERES:\n(0, 0, 3), 1, 10\n",
        "bad_float.scene": "MATERIALS:\ndiffuse, (1, 1, 1), 1e3\n",
        "missing_param.scene": "MATERIALS:\ndiffuse, (1, 1, 1), 1\nSPHERES:\n(0, 0, 3), 1\n",
        "no_section.scene": "(0, 0, 3), 1, 0\n",
        "bad_operation.scene": "MATERIALS:\ndiffuse, (1, 1, 1), 1\nrotate: 3, (0, 1, 0)\n",
        "blank_line.scene": "MATERIALS:\ndiffuse, (1, 1, 1), 1\n   \n",
        "bad_path.scene": "MATERIALS:\ntextured, (1, 1, 1), 1\nMODELS:\nload: shapes.obj, 0\n",
        "no_material.scene": "MATERIALS:\ndiffuse, (1, 1, 1), 1\nMODELS:\nload: \"shapes.obj\", 3\n",
        "untextured_model.scene": "MATERIALS:\ndiffuse, (1, 1, 1), 1\nMODELS:\nload: \"shapes.obj\", 0\n",
    }
    for name, text in bad.items():
        with open(os.path.join(HERE, name), "w") as f:
            f.write(text)


if __name__ == "__main__":
    main()
