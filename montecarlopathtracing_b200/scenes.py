"""The three course scenes (Scene/README.md:20-61; config.json:3-84) prepared through the host layer."""
import os
from dataclasses import dataclass

import numpy as np

from . import host

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ASSETS = os.path.join(ROOT, "assets")

SCENES = {
    # name: (directory, objname, default resolution, default maxdepth)
    "cbox": ("cbox", "cbox.obj", (512, 512), 16),
    "veach_mis": ("veach_mis", "mis.obj", (768, 512), 16),
    "diningroom": ("diningroom", "diningroom.obj", (1280, 720), 8),
}


@dataclass
class PreparedScene:
    name: str
    triangles: np.ndarray
    materials: np.ndarray
    nodes: np.ndarray
    camera: np.ndarray

    @property
    def scene_bytes(self):
        return self.triangles.nbytes + self.materials.nbytes + self.nodes.nbytes + self.camera.nbytes


def scene_dir(name):
    return os.path.join(ASSETS, SCENES[name][0]) + os.sep


def load(name, bvhtype="hlbvh"):
    d, obj, _, _ = SCENES[name]
    path = os.path.join(ASSETS, d, obj)
    if not os.path.exists(path):
        raise FileNotFoundError(path + " (diningroom is a procedural stand-in: run tools/make_diningroom.py)")
    sc = host.load_object(scene_dir(name), obj)
    nodes = host.build_bvh(sc.triangles, bvhtype)
    cam = host.parse_camera(host.CAMERAS[name])
    return PreparedScene(name, sc.triangles, sc.materials, nodes, cam)
