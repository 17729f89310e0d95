"""The host loader (ray-tracer-challenge_b200/csrc/host: JSON -> World tree -> blob) checked against a restatement
that shares NO code with it: ~100 lines of Python / numpy following parseObject's inherit -> build -> setTransform
push-down -> material / casts-shadow stamping order (/root/reference/src/parsing/scene.zig:164-190, 214-241, 407-430,
440-591; shapes/shape.zig:286-310), compared with what the blob holds for every leaf: world->object matrix, material
numbers, casts_shadow, kind.  Why: the oracle links the product's loader, so GPU-vs-oracle parity
cannot see a loader deviation; this test (and the transcribed scene.zig / obj.zig KATs in oracle/kat_tests.cpp) can.
divide(8) only regroups leaves, so the comparison is between multisets of leaves."""
import collections
import json
import os
import struct

import numpy as np
import pytest

from tests import scenes

DEFAULT_MATERIAL = dict(ambient=0.1, diffuse=0.9, specular=0.9, shininess=200.0, reflective=0.0, transparency=0.0,
                        refractive_index=1.0)  # material.zig:27-38


def parse_transform(ops):  # scene.zig:214-241: every op multiplies from the LEFT (matrix.zig:222-325)
    m = np.eye(4)
    for op in ops:
        (name, v), = op.items()
        t = np.eye(4)
        if name == "translate":
            t[:3, 3] = v
        elif name == "scale":
            t[0, 0], t[1, 1], t[2, 2] = v
        elif name in ("rotate-x", "rotate-y", "rotate-z"):
            c, s = np.cos(v), np.sin(v)
            a, b = {"rotate-x": (1, 2), "rotate-y": (2, 0), "rotate-z": (0, 1)}[name]
            t[a, a], t[a, b], t[b, a], t[b, b] = c, -s, s, c
        elif name == "shear":
            t[0, 1], t[0, 2], t[1, 0], t[1, 2], t[2, 0], t[2, 1] = (v[k] for k in ("xy", "xz", "yx", "yz", "zx", "zy"))
        else:
            raise ValueError(name)
        m = t @ m
    return m


def parse_material(cfg, inherited):  # scene.zig:407-430 (the pattern is not compared here)
    mat = dict(inherited if inherited is not None else DEFAULT_MATERIAL)
    for k in ("ambient", "diffuse", "specular", "shininess", "reflective", "transparency"):
        if cfg.get(k) is not None:
            mat[k] = float(cfg[k])
    if cfg.get("refractive-index") is not None:
        mat["refractive_index"] = float(cfg["refractive-index"])
    return mat


def inherit(obj, state):  # scene.zig:164-190
    material = parse_material(obj["material"], state["material"]) if obj.get("material") is not None else state["material"]
    transform = parse_transform(obj["transform"]) @ state["transform"] if obj.get("transform") is not None else state["transform"]
    shadow = obj["casts-shadow"] if obj.get("casts-shadow") is not None else state["casts_shadow"]
    return dict(material=material, transform=transform, casts_shadow=shadow)


class Node:
    def __init__(self, kind, children=None, n=1):
        self.kind, self.children, self.n = kind, children, n  # n: identical leaves this node stands for (OBJ triangles)
        self.transform = np.eye(4)  # Shape._transform; groups and CSGs never set their own (shape.zig:286-302)
        self.material, self.casts_shadow = dict(DEFAULT_MATERIAL), True

    def set_transform(self, m):  # shape.zig:286-310
        if self.children is not None:
            for c in self.children:
                c.set_transform(m @ c.transform)
        else:
            self.transform = m


def obj_triangles(path):  # obj.zig:195-279: a face of k vertices is a fan of k - 2 triangles
    n = 0
    with open(path) as f:
        for line in f:
            p = line.split()
            if p and p[0] == "f":
                n += max(0, len(p) - 3)
    return n


def parse_object(obj, state, defs, asset_dir):  # scene.zig:440-591
    info = inherit(obj, state)
    material, transform, shadow = info["material"], info["transform"], info["casts_shadow"]
    typ = obj["type"]
    (name, arg), = typ.items() if isinstance(typ, dict) else ((typ, None),)
    if name == "from-definition":
        parent = parse_object(defs[arg], dict(material=material, transform=state["transform"], casts_shadow=shadow), defs, asset_dir)
        new = inherit(obj, dict(material=parent.material, transform=parent.transform, casts_shadow=parent.casts_shadow))
        material, transform, shadow = new["material"], new["transform"], new["casts_shadow"]
        shape = parent
    elif name == "from-obj":
        tri = Node("triangle", n=obj_triangles(os.path.join(asset_dir, arg["file"])))
        tri.material = dict(material) if material is not None else dict(DEFAULT_MATERIAL)  # stamped by loadObj
        tri.casts_shadow = shadow if shadow is not None else True
        shape = Node("group", [tri])
    elif name == "group":
        shape = Node("group", [parse_object(c, dict(material=material, transform=np.eye(4), casts_shadow=shadow), defs, asset_dir)
                               for c in arg])
    elif name == "csg":
        kids = [parse_object(arg[k], dict(material=material, transform=np.eye(4), casts_shadow=shadow), defs, asset_dir)
                for k in ("left", "right")]
        shape = Node("csg", kids)
    else:
        shape = Node(name)
    shape.set_transform(transform)
    if material is not None:
        shape.material = dict(material)
    if shadow is not None:
        shape.casts_shadow = shadow
    return shape


def leaves_of(node, out):
    if node.children is not None:
        for c in node.children:
            leaves_of(c, out)
    else:
        out.append(node)


def key_of(kind, inv12, mat, shadow):
    scale = max(1e-300, max(abs(float(v)) for v in inv12))  # 9 digits relative to the matrix's largest entry (noise near 0)
    m = tuple(round(float(v) / scale, 9) * scale + 0.0 for v in inv12)
    m = tuple(float("%.8e" % v) + 0.0 for v in m)
    return (kind, m, tuple(float("%.12e" % mat[k]) for k in sorted(DEFAULT_MATERIAL)), bool(shadow))


def blob_leaves(blob):
    """(kind, inverse rows 0..2, material numbers, casts_shadow) of every leaf, read straight from the blob (rtc_blob.h)."""
    counts = struct.unpack_from("<12I", blob, 16)
    offs = struct.unpack_from("<14Q", blob, 16 + 48 + 8)
    n_mat, n_xf, n_leaves = counts[1], counts[5], counts[6]
    off_mat, off_xf, off_leaves = offs[1], offs[5], offs[6]
    mats = [struct.unpack_from("<7d", blob, off_mat + 64 * i) for i in range(n_mat)]
    xfs = np.frombuffer(blob, dtype="<f8", count=12 * n_xf, offset=off_xf).reshape(n_xf, 12)
    names = ["sphere", "plane", "cube", "cylinder", "cone", "triangle", "triangle", "test"]
    out = collections.Counter()
    for i in range(n_leaves):
        kind, xf, mat, flags = struct.unpack_from("<iiiI", blob, off_leaves + 96 * i)
        m = dict(zip(("ambient", "diffuse", "specular", "shininess", "reflective", "transparency", "refractive_index"), mats[mat]))
        out[key_of(names[kind], xfs[xf], m, flags & 1)] += 1
    return out


@pytest.mark.parametrize("name", ["dragons", "groups", "cover", "csg", "cylinders", "teapot"])
def test_blob_leaves_match_an_independent_restatement_of_parse_object(pkg, name):
    s = scenes.shipped_scene(name)
    if s is None:
        pytest.skip("reference assets not staged (assets/_ref)")
    text, assets = s
    cfg = json.loads(text)
    defs = {d["name"]: d["value"] for d in cfg.get("shape-definitions", [])}
    want = collections.Counter()
    for obj in cfg["objects"]:
        root = parse_object(obj, dict(material=None, transform=np.eye(4), casts_shadow=None), defs, assets)
        leaves = []
        leaves_of(root, leaves)
        for lf in leaves:
            inv = np.linalg.inv(lf.transform)
            want[key_of(lf.kind, inv[:3].reshape(12), lf.material, lf.casts_shadow)] += lf.n
    blob, _ = pkg.load_blob(text, assets)
    got = blob_leaves(blob)
    assert sum(got.values()) == sum(want.values())
    missing = {k: v for k, v in want.items() if got.get(k) != v}
    assert not missing, "leaves the blob does not hold as the reference's parse order gives them: %r" % list(missing.items())[:3]
