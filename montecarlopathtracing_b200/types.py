"""numpy views of the POD records in include/mcpt_types.h (byte-identical to the reference's
objdef.h:21-99; sizes asserted below and in tests/test_host.py)."""
import numpy as np

F4 = (np.float32, (4,))

CAMERA = np.dtype([("center", *F4), ("direction", *F4), ("up", *F4), ("horizontal", *F4),
                   ("arg", np.float32), ("tmin", np.float32), ("camera_type", np.uint32),
                   ("pad_", np.float32)])
RAY = np.dtype([("origin", np.float32, (3,)), ("state", np.int32),
                ("direction", np.float32, (3,)), ("pixel", np.int32), ("ratio", *F4)])
HIT = np.dtype([("normal", *F4), ("point", *F4), ("t", np.float32), ("triangle_id", np.uint32),
                ("material_id", np.uint32), ("pad_", np.uint32)])
TRIANGLE = np.dtype([("v", np.float32, (3, 4)), ("normal", np.float32, (3,)), ("material_id", np.int32)])
MATERIAL = np.dtype([("type", np.int32), ("Ni", np.float32), ("Ns", np.float32), ("pad_", np.float32),
                     ("kd", *F4), ("ka_ks", *F4)])
BVH_NODE = np.dtype([("bbmin", *F4), ("bbmax", *F4), ("pad_", *F4), ("parent", np.int32),
                     ("left", np.int32), ("right", np.int32), ("pad2_", np.int32)])

assert CAMERA.itemsize == 80 and RAY.itemsize == 48 and HIT.itemsize == 48
assert TRIANGLE.itemsize == 64 and MATERIAL.itemsize == 48 and BVH_NODE.itemsize == 64

DIFFUSE, GLOSSY, TRANSPARENT, LIGHT = 1, 2, 3, 4
RAY_DEPTH_MASK, RAY_INSIDE_MASK, RAY_TERM_MASK = 0x0000FFFF, 0x00FF0000, 0xFF000000
