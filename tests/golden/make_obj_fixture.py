#!/usr/bin/env python
"""Writes tests/golden/own_box.obj + own_box.mtl: a small OBJ/MTL scene of OUR OWN content (no reference data) that
exercises every branch of csrc/obj_loader.cpp on machines without /root/reference (the GPU box): CRLF line ends, quads and a
pentagon (fan triangulation), v/vt/vn, v//vn, v/vt and bare v corners, negative indices, faces before any usemtl, comments,
o/g/s lines, and one MTL entry per branch of the reference's Ks = (roughness, metalness, +-ior) decode (Renderer.mm:292-320).
Deterministic; rerun = identical bytes."""
import pathlib

here = pathlib.Path(__file__).resolve().parent
L = []
def v(*p): L.append("v %g %g %g" % p)
def vn(*p): L.append("vn %g %g %g" % p)

L += ["# own_box.obj - written by make_obj_fixture.py", "mtllib own_box.mtl", "o room"]
# room: x in [-1,1], y in [0,2], z in [-1,1]; open towards +z (the camera sits at z = 2.35)
for p in [(-1, 0, -1), (1, 0, -1), (1, 0, 1), (-1, 0, 1), (-1, 2, -1), (1, 2, -1), (1, 2, 1), (-1, 2, 1)]:
    v(*p)
for n in [(0, 1, 0), (0, -1, 0), (0, 0, 1), (1, 0, 0), (-1, 0, 0)]:
    vn(*n)
L += ["vt 0 0", "vt 1 0", "vt 1 1", "vt 0 1"]
L += ["f 1//1 4//1 3//1 2//1"]                       # floor, before any usemtl -> "__default__" material (white diffuse)
L += ["usemtl white", "s off", "f 5/1/2 6/2/2 7/3/2 8/4/2", "f 1//3 2//3 6//3 5//3"]       # ceiling (v/vt/vn), back wall
L += ["usemtl red", "f 1/1 5/2 8/3 4/4"]             # left wall: v/vt corners, no normals -> face normal
L += ["usemtl green", "f -7 -6 -2 -3"]               # right wall by negative indices (2 3 7 6), bare v corners
L += ["g light", "usemtl light"]
for p in [(-0.3, 1.98, -0.3), (0.3, 1.98, -0.3), (0.3, 1.98, 0.3), (-0.3, 1.98, 0.3)]:
    v(*p)
L += ["f 9//2 10//2 11//2 12//2"]
# glass block (closed box, outward normals from the faces), x [-0.6,-0.1], y [0,0.7], z [-0.3,0.3]
L += ["o glass", "usemtl glass"]
def box(x0, y0, z0, x1, y1, z1, first):
    for p in [(x0, y0, z0), (x1, y0, z0), (x1, y0, z1), (x0, y0, z1), (x0, y1, z0), (x1, y1, z0), (x1, y1, z1), (x0, y1, z1)]:
        v(*p)
    a = first
    for q in [(0, 1, 2, 3), (4, 7, 6, 5), (0, 4, 5, 1), (1, 5, 6, 2), (2, 6, 7, 3), (3, 7, 4, 0)]:
        L.append("f " + " ".join(str(a + k) for k in q))
box(-0.6, 0.0, -0.3, -0.1, 0.7, 0.3, 13)
L += ["usemtl plastic"]; box(0.2, 0.0, -0.5, 0.7, 0.4, 0.0, 21)
L += ["usemtl mirror"]
for p in [(-0.9, 0.3, -0.99), (0.9, 0.3, -0.99), (0.9, 1.2, -0.99), (0.0, 1.6, -0.99), (-0.9, 1.2, -0.99)]:
    v(*p)
L += ["f 29//3 30//3 31//3 32//3 33//3"]             # pentagon -> 3 triangles
L += ["usemtl brushed", "f 21 22 26"]                # rough conductor: stays diffuse; a second use of plastic's vertices
L += ["usemtl frosted", "f 22 23 27"]                # rough dielectric: diffuse
L += ["usemtl white", "f 25 26 27"]                  # back to an earlier material: joins its group
L += ["usemtl missing", "f 13 14 18", ""]            # not in the MTL: default entry (white diffuse)
(here / "own_box.obj").write_bytes("\r\n".join(L).encode())

M = ["# own_box.mtl", "newmtl white", "Kd 0.73 0.73 0.73", "Ks 1 0 0", "",
     "newmtl red", "Kd 0.65 0.05 0.05", "Ks 1 0 1.2",
     "newmtl green", "Kd 0.12 0.45 0.15",
     "newmtl light", "Kd 0.8 0.8 0.8", "Ka 10 9 8",
     "newmtl glass", "Kd 1 1 1", "Ks 0 0 1.5",
     "newmtl plastic", "Kd 0.3 0.55 0.85", "Ks 0 0 -1.5",
     "newmtl mirror", "Kd 0.95 0.95 0.95", "Ks 0 1 0",
     "newmtl brushed", "Kd 0.6 0.6 0.6", "Ks 0.5 1 0",
     "newmtl frosted", "Kd 0.7 0.7 0.7", "Ks 0.3 0 1.5", ""]
(here / "own_box.mtl").write_bytes("\r\n".join(M).encode())
print("wrote own_box.obj (%d lines), own_box.mtl" % len(L))
