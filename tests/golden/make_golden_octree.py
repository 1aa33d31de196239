"""Generates tests/golden/octree_*.npz from the REFERENCE's own oct-tree builder (RayTracer::createOctTree / loadOctTree /
refitOctTreeNode, RayTracer.cpp:503-788, compiled by oracle/Makefile into oracle/_ref/libokrt_ref_octree.so).  Run HERE, where
/root/reference exists:   python tests/golden/make_golden_octree.py
Each fixture holds the inputs (mesh boxes, mesh / sphere entity arrays, maxDepth, maxLeafObjects) and the reference's outputs
(GPU_OctTreeNode[], GPU_MeshComponent[]).  The meshes are reduced to one triangle spanning their bounding box: createOctTree
reads nothing of a mesh but its box and its MeshDesc."""
import ctypes
import importlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
HERE = os.path.dirname(os.path.abspath(__file__))


def box_meshes(okrt, boxes):
    meshes = []
    for lo, hi in boxes:
        tri = np.zeros(1, okrt.abi.TRIANGLE)
        tri[0]["position"] = [lo, hi, [lo[0], hi[1], lo[2]]]
        meshes.append(okrt.Mesh(tri, None, (np.asarray(lo, np.float32), np.asarray(hi, np.float32))))
    return meshes


class EntityArrays:
    """stands in for okrt.Scene where only the entity arrays are needed"""
    def __init__(self, me, se):
        self._me, self._se = me, se

    def meshEntityArray(self):
        return self._me

    def sphereEntityArray(self):
        return self._se


def main():
    okrt = importlib.import_module("gpu-raytracer_b200")
    abi = okrt.abi
    from oracle import binding
    ref = binding.load_ref_octree()
    assert ref is not None, "needs /root/reference (oracle/_ref/libokrt_ref_octree.so)"
    rng = np.random.default_rng(23)
    rm2, sc2, _ = okrt.scenes.moved_props_scene()
    for k in range(6):
        sc2.addSphereEntity(float(rng.uniform(0.5, 3.0)), okrt.scenes.material(), tuple(rng.uniform(-20, 20, 3)))
    cases = {"testscene_d3_l2": (okrt.scenes.test_scene()[:2], 3, 2), "moved_props_spheres_d3_l3": ((rm2, sc2), 3, 3),
             "sponza8k_d2_l50": (okrt.scenes.sponza_standin(8000, texture_size=16)[:2], 2, 50)}
    for name, ((rm, sc), depth, leaf) in cases.items():
        boxes = np.array([[m.boundingBox[0], m.boundingBox[1]] for m in rm.meshes], np.float32)
        geo = okrt.Geometry(box_meshes(okrt, boxes), 30, 5)
        me, se = sc.meshEntityArray(), sc.sphereEntityArray()
        nodes = np.zeros(8192, abi.OCTREE_NODE); inst = np.zeros(max(1, len(me)), abi.MESH_INSTANCE); ni = ctypes.c_uint32()
        bounds = np.ascontiguousarray(geo.meshBounds)
        nn = ref.ref_build_octree(abi.ptr(geo.meshDescs), abi.ptr(bounds), len(geo.meshDescs), abi.ptr(me), len(me),
                                  abi.ptr(se) if len(se) else None, len(se), depth, leaf, abi.ptr(nodes), len(nodes), abi.ptr(inst), len(inst), ctypes.byref(ni))
        assert nn != 0xFFFFFFFF
        np.savez_compressed(os.path.join(HERE, "octree_%s.npz" % name), boxes=boxes, meshEntities=me.view(np.uint8), sphereEntities=se.view(np.uint8),
                            maxDepth=depth, maxLeafObjects=leaf, nodes=nodes[:nn].view(np.uint8), instances=inst[:ni.value].view(np.uint8))
        print("octree", name, len(me), "mesh entities", len(se), "spheres ->", nn, "nodes")


if __name__ == "__main__":
    main()
