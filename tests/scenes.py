"""Scenes written for the tests (own JSON in the reference's schema, src/parsing/scene.zig:28-210),
so that parity tests do not depend on the staged reference assets."""
import json
import math
import os
import random

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ASSETS = os.path.join(ROOT, "assets", "_ref")
SHIPPED = ["align_check", "cover", "csg", "cubes", "cylinders", "dragons", "earth", "fresnel", "groups", "nefertiti",
           "reflection_and_refraction", "skybox", "teapot", "xyz"]


def shipped_scene(name):
    """-> (json text, asset dir) of a scene staged from the reference mount, or None."""
    p = os.path.join(ASSETS, "scenes", name + ".json")
    if not os.path.exists(p):
        return None
    with open(p) as f:
        return f.read(), os.path.join(ASSETS, "data")


def solid(r, g, b):
    return {"type": {"solid": [r, g, b]}}


def camera(w, h, fov=0.9, frm=(0, 2.5, -7), to=(0, 0.8, 0), up=(0, 1, 0)):
    return {"width": w, "height": h, "field-of-view": fov, "from": list(frm), "to": list(to), "up": list(up)}


def light(pos, inten=(1, 1, 1)):
    return {"point-light": {"position": list(pos), "intensity": list(inten)}}


def _smoke():
    objs = [
        {"type": {"plane": {}},
         "material": {"pattern": {"type": {"checkers": [solid(0.2, 0.2, 0.2), solid(0.9, 0.9, 0.9)]},
                                  "transform": [{"scale": [0.7, 0.7, 0.7]}]},
                      "reflective": 0.3, "specular": 0.1}},
        {"type": {"sphere": {}}, "transform": [{"translate": [-1.6, 1, 0.3]}],
         "material": {"pattern": solid(1, 1, 1), "ambient": 0, "diffuse": 0.05, "specular": 0.9, "shininess": 300,
                      "reflective": 0.9, "transparency": 0.9, "refractive-index": 1.5}},
        {"type": {"sphere": {}}, "transform": [{"scale": [0.5, 0.5, 0.5]}, {"translate": [-1.6, 1, 0.3]}],
         "material": {"pattern": solid(1, 1, 1), "ambient": 0, "diffuse": 0, "transparency": 1.0,
                      "reflective": 0.5, "refractive-index": 1.0003}},
        {"type": {"cube": {}}, "transform": [{"rotate-y": 0.6}, {"scale": [0.6, 0.6, 0.6]}, {"translate": [1.5, 0.6, -0.5]}],
         "material": {"pattern": {"type": {"stripes": [solid(0.9, 0.2, 0.1), solid(0.1, 0.2, 0.9)]},
                                  "transform": [{"scale": [0.25, 0.25, 0.25]}, {"rotate-z": 0.5}]}}},
        {"type": {"cylinder": {"min": 0, "max": 1.5, "closed": True}},
         "transform": [{"scale": [0.4, 1, 0.4]}, {"translate": [0.2, 0, 1.8]}],
         "material": {"pattern": {"type": {"gradient": [solid(0.1, 0.8, 0.2), solid(0.9, 0.9, 0.1)]},
                                  "transform": [{"rotate-z": 1.5708}, {"scale": [1, 1.5, 1]}]},
                      "reflective": 0.2}},
        {"type": {"cone": {"min": -1, "max": 0, "closed": True}},
         "transform": [{"translate": [0, 1, 0]}, {"scale": [0.6, 1.2, 0.6]}, {"translate": [3.0, 0, 1.5]}],
         "material": {"pattern": {"type": {"rings": [solid(0.8, 0.5, 0.1), solid(0.3, 0.1, 0.5)]},
                                  "transform": [{"scale": [0.15, 0.15, 0.15]}]}}},
        {"type": {"csg": {"operation": "difference",
                          "left": {"type": {"cube": {}}, "material": {"pattern": solid(0.9, 0.8, 0.1)}},
                          "right": {"type": {"sphere": {}}, "transform": [{"scale": [1.3, 1.3, 1.3]}],
                                    "material": {"pattern": solid(0.8, 0.1, 0.1)}}}},
         "transform": [{"scale": [0.5, 0.5, 0.5]}, {"rotate-y": 0.4}, {"translate": [-0.2, 0.5, -2.2]}]},
        {"type": {"group": [
            {"type": {"sphere": {}}, "transform": [{"scale": [0.3, 0.3, 0.3]}, {"translate": [0, 0.3, 0]}]},
            {"type": {"sphere": {}}, "transform": [{"scale": [0.3, 0.3, 0.3]}, {"translate": [0.7, 0.3, 0]}],
             "casts-shadow": False},
            {"type": {"triangle": {"p1": [0, 0, 0], "p2": [0.7, 0, 0], "p3": [0.35, 1.2, 0.3]}}}]},
         "transform": [{"rotate-y": -0.5}, {"translate": [-3.2, 0, -1.5]}],
         "material": {"pattern": solid(0.2, 0.7, 0.7), "reflective": 0.1}},
    ]
    return {"camera": camera(96, 64), "lights": [light((-5, 8, -6), (0.8, 0.8, 0.8)), light((6, 5, -3), (0.35, 0.3, 0.3))],
            "objects": objs}


SMOKE_SCENE = json.dumps(_smoke())


def default_world(w=11, h=11):
    """World.default (world.zig:40-62) seen by the camera of camera.zig:171-187."""
    return json.dumps({
        "camera": camera(w, h, math.pi / 2, (0, 0, -5), (0, 0, 0), (0, 1, 0)),
        "lights": [light((-10, 10, -10))],
        "objects": [
            {"type": {"sphere": {}}, "material": {"pattern": solid(0.8, 1.0, 0.6), "diffuse": 0.7, "specular": 0.2}},
            {"type": {"sphere": {}}, "transform": [{"scale": [0.5, 0.5, 0.5]}]}]})


def write_mesh_obj(path, n_u=24, n_v=16, smooth=True, seed=7, groups=False):
    """A bumpy torus as OBJ (v / vn / f with quads -> fan triangulation, optional `g` groups)."""
    rnd = random.Random(seed)
    lines, R, r = [], 1.0, 0.42
    for i in range(n_u):
        for j in range(n_v):
            a, b = 2 * math.pi * i / n_u, 2 * math.pi * j / n_v
            rr = r * (1 + 0.15 * math.sin(3 * a) * math.cos(2 * b)) + 0.01 * rnd.random()
            x, y, z = (R + rr * math.cos(b)) * math.cos(a), rr * math.sin(b), (R + rr * math.cos(b)) * math.sin(a)
            lines.append("v %.9f %.9f %.9f" % (x, y, z))
    if smooth:
        for i in range(n_u):
            for j in range(n_v):
                a, b = 2 * math.pi * i / n_u, 2 * math.pi * j / n_v
                lines.append("vn %.9f %.9f %.9f" % (math.cos(b) * math.cos(a), math.sin(b), math.cos(b) * math.sin(a)))
    for i in range(n_u):
        if groups and i % 8 == 0:
            lines.append("g part%d" % (i // 8))
        for j in range(n_v):
            q = [i * n_v + j, ((i + 1) % n_u) * n_v + j, ((i + 1) % n_u) * n_v + (j + 1) % n_v, i * n_v + (j + 1) % n_v]
            if smooth:
                lines.append("f " + " ".join("%d//%d" % (k + 1, k + 1) for k in q))
            else:
                lines.append("f " + " ".join("%d" % (k + 1) for k in q))
    with open(path, "w") as f:
        f.write("\n".join(lines) + "\n")


def mesh_scene(obj_name, w=120, h=80, normalize=True):
    """Two instances of one OBJ (one refractive), a mirror floor, two lights."""
    return json.dumps({
        "shape-definitions": [{"name": "torus", "value": {"type": {"from-obj": {"file": obj_name, "normalize": normalize}}}}],
        "camera": camera(w, h, 0.8, (0, 3.0, -6), (0, 0.6, 0)),
        "lights": [light((-4, 7, -5), (0.9, 0.9, 0.9)), light((5, 4, -2), (0.3, 0.3, 0.4))],
        "objects": [
            {"type": {"plane": {}}, "material": {"pattern": {"type": {"checkers": [solid(0.1, 0.1, 0.1), solid(0.8, 0.8, 0.8)]}},
                                                  "reflective": 0.25}},
            {"type": {"from-definition": "torus"}, "transform": [{"rotate-x": 0.5}, {"translate": [-1.3, 1.0, 0]}],
             "material": {"pattern": solid(0.8, 0.3, 0.2), "reflective": 0.2}},
            {"type": {"from-definition": "torus"},
             "transform": [{"scale": [0.8, 0.8, 0.8]}, {"rotate-z": 0.9}, {"translate": [1.4, 0.9, 0.5]}],
             "material": {"pattern": solid(0.9, 0.9, 1.0), "diffuse": 0.1, "ambient": 0.05, "transparency": 0.8,
                          "reflective": 0.3, "refractive-index": 1.3}},
            {"type": {"cube": {}}, "transform": [{"scale": [0.5, 0.5, 0.5]}, {"translate": [0, 0.5, 2.5]}],
             "material": {"pattern": solid(0.2, 0.5, 0.9), "transparency": 0.6, "refractive-index": 1.0},
             "casts-shadow": False},
        ]})


def with_size(json_text, w, h):
    d = json.loads(json_text)
    d["camera"]["width"], d["camera"]["height"] = w, h
    return json.dumps(d)
