#!/usr/bin/env python
"""Generates the wire-format fixtures under tests/golden/:

  wire_hybrid_grid.bin        serialized cartographer.mapping.proto.HybridGrid
  wire_hybrid_grid_tsdf.bin   serialized cartographer.mapping.proto.HybridGridTSDF
  wire_unpacked.bin           the HybridGrid again with [packed = false] repeated fields
  wire_map.pbstream           a proto stream: header, a pose-graph stand-in, two Submap3D, a node
  wire_expected.npz           the voxels / submap fields the files hold

The messages are ENCODED BY GOOGLE'S PROTOBUF RUNTIME (python package, descriptors built at run
time from the field numbers and types of the reference's .proto files — there is no protoc in
this image), so the library's own decoder is checked against an independent encoder. The .proto
sources restated below: cartographer/mapping/proto/3d/hybrid_grid.proto:19-29,
3d/hybrid_grid_tsdf.proto:19-32, submap.proto:33-47, serialization.proto:26-30 and 72-88,
pose_graph.proto:22-25, transform/proto/transform.proto:33-75; container: io/proto_stream.cc:27-60.
Regenerate with:  python tests/golden/make_wire_golden.py
"""
import gzip
import os
import struct
import sys

import numpy as np
from google.protobuf import descriptor_pb2, descriptor_pool, message_factory

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

T = descriptor_pb2.FieldDescriptorProto
MAGIC = 0x7b1d1f7b5bf501db


def _message(fd, name, fields):
    m = fd.message_type.add()
    m.name = name
    for fname, number, ftype, label, type_name, packed in fields:
        f = m.field.add()
        f.name, f.number, f.type, f.label = fname, number, ftype, label
        if type_name:
            f.type_name = type_name
        if packed is not None:
            f.options.packed = packed
    return m


def build_messages():
    OPT, REP = T.LABEL_OPTIONAL, T.LABEL_REPEATED
    fd = descriptor_pb2.FileDescriptorProto()
    fd.name = "hg_wire_golden.proto"
    fd.package = "hgw"
    fd.syntax = "proto3"
    grid_fields = [("resolution", 1, T.TYPE_FLOAT, OPT, None, None),
                   ("x_indices", 3, T.TYPE_SINT32, REP, None, None),
                   ("y_indices", 4, T.TYPE_SINT32, REP, None, None),
                   ("z_indices", 5, T.TYPE_SINT32, REP, None, None)]
    _message(fd, "HybridGrid", grid_fields + [("values", 6, T.TYPE_INT32, REP, None, None)])
    _message(fd, "HybridGridUnpacked",
             [(n, k, t, l, tn, (False if l == REP else None)) for n, k, t, l, tn, _ in
              grid_fields + [("values", 6, T.TYPE_INT32, REP, None, None)]])
    _message(fd, "HybridGridTSDF", grid_fields + [
        ("values_tsd", 6, T.TYPE_INT32, REP, None, None), ("values_weight", 7, T.TYPE_INT32, REP, None, None),
        ("relative_truncation_distance", 8, T.TYPE_FLOAT, OPT, None, None),
        ("max_weight", 9, T.TYPE_FLOAT, OPT, None, None)])
    _message(fd, "Vector3d", [(n, i + 1, T.TYPE_DOUBLE, OPT, None, None) for i, n in enumerate("xyz")])
    _message(fd, "Quaterniond", [(n, i + 1, T.TYPE_DOUBLE, OPT, None, None) for i, n in enumerate("xyzw")])
    _message(fd, "Rigid3d", [("translation", 1, T.TYPE_MESSAGE, OPT, ".hgw.Vector3d", None),
                             ("rotation", 2, T.TYPE_MESSAGE, OPT, ".hgw.Quaterniond", None)])
    _message(fd, "Submap3D", [
        ("local_pose", 1, T.TYPE_MESSAGE, OPT, ".hgw.Rigid3d", None),
        ("num_range_data", 2, T.TYPE_INT32, OPT, None, None), ("finished", 3, T.TYPE_BOOL, OPT, None, None),
        ("high_resolution_hybrid_grid", 4, T.TYPE_MESSAGE, OPT, ".hgw.HybridGrid", None),
        ("low_resolution_hybrid_grid", 5, T.TYPE_MESSAGE, OPT, ".hgw.HybridGrid", None),
        ("rotational_scan_matcher_histogram", 6, T.TYPE_FLOAT, REP, None, None),
        ("high_resolution_hybrid_grid_tsdf", 7, T.TYPE_MESSAGE, OPT, ".hgw.HybridGridTSDF", None),
        ("low_resolution_hybrid_grid_tsdf", 8, T.TYPE_MESSAGE, OPT, ".hgw.HybridGridTSDF", None)])
    _message(fd, "SubmapId", [("trajectory_id", 1, T.TYPE_INT32, OPT, None, None),
                              ("submap_index", 2, T.TYPE_INT32, OPT, None, None)])
    _message(fd, "Submap", [("submap_id", 1, T.TYPE_MESSAGE, OPT, ".hgw.SubmapId", None),
                            ("submap_3d", 3, T.TYPE_MESSAGE, OPT, ".hgw.Submap3D", None)])
    _message(fd, "Opaque", [("payload", 1, T.TYPE_BYTES, OPT, None, None),
                            ("number", 2, T.TYPE_INT64, OPT, None, None)])
    _message(fd, "SerializationHeader", [("format_version", 1, T.TYPE_UINT32, OPT, None, None)])
    _message(fd, "SerializedData", [("pose_graph", 1, T.TYPE_MESSAGE, OPT, ".hgw.Opaque", None),
                                    ("submap", 3, T.TYPE_MESSAGE, OPT, ".hgw.Submap", None),
                                    ("node", 4, T.TYPE_MESSAGE, OPT, ".hgw.Opaque", None)])
    pool = descriptor_pool.DescriptorPool()
    pool.Add(fd)
    return {m.name: message_factory.GetMessageClass(pool.FindMessageTypeByName("hgw." + m.name))
            for m in fd.message_type}


def fill_grid(msg, xyz, a, b=None, resolution=0.1, rtd=None, max_weight=None):
    msg.resolution = resolution
    msg.x_indices.extend(int(v) for v in xyz[:, 0])
    msg.y_indices.extend(int(v) for v in xyz[:, 1])
    msg.z_indices.extend(int(v) for v in xyz[:, 2])
    if b is None:
        msg.values.extend(int(v) for v in a)
    else:
        msg.values_tsd.extend(int(v) for v in a)
        msg.values_weight.extend(int(v) for v in b)
        msg.relative_truncation_distance = rtd
        msg.max_weight = max_weight
    return msg


def main():
    from hectorgrapher_b200 import synth
    from oracle import pyoracle
    pyoracle.build()
    M = build_messages()
    sph = synth.spheres()
    origin, yaw = synth.corkscrew_pose(0, 2)
    cloud = synth.scan(origin, yaw, 8, 90, sph, np.random.RandomState(9))
    world = (synth.quat_rotate(synth.quat_from_yaw(yaw), cloud.astype(np.float64)) + origin).astype(np.float32)
    prob = pyoracle.Grid.prob(0.10)
    prob.insert_occupancy(origin.astype(np.float32), world, 0.55, 0.49, 2)
    prob_lo = pyoracle.Grid.prob(0.45)
    prob_lo.insert_occupancy(origin.astype(np.float32), world, 0.55, 0.49, 2)
    tsdf = pyoracle.Grid.tsdf(0.10, 2.5, 1000.0)
    tsdf.insert_tsdf(origin.astype(np.float32), world, epsilon=1.0)
    tsdf_lo = pyoracle.Grid.tsdf(0.45, 2.5, 1000.0)
    tsdf_lo.insert_tsdf(origin.astype(np.float32), world, epsilon=1.0)
    exp = {}
    px, pa = prob.export()
    # a cell far out on the negative side exercises zig-zag and multi-byte varints
    px = np.concatenate([px, np.array([[-8000, 8000, -1]], np.int32)])
    pa = np.concatenate([pa, np.array([32767], np.uint16)])
    exp["prob_xyz"], exp["prob_values"] = px, pa
    open(os.path.join(HERE, "wire_hybrid_grid.bin"), "wb").write(
        fill_grid(M["HybridGrid"](), px, pa).SerializeToString())
    open(os.path.join(HERE, "wire_unpacked.bin"), "wb").write(
        fill_grid(M["HybridGridUnpacked"](), px, pa).SerializeToString())
    tx, ta, tb = tsdf.export()
    exp["tsdf_xyz"], exp["tsdf_tsd"], exp["tsdf_weight"] = tx, ta, tb
    # HybridGridTSDF::ToProto stores getMaxTSD() (= 2.5 * 0.1) in the relative_truncation_distance field
    open(os.path.join(HERE, "wire_hybrid_grid_tsdf.bin"), "wb").write(
        fill_grid(M["HybridGridTSDF"](), tx, ta, tb, 0.1, float(np.float32(2.5) * np.float32(0.1)), 1000.0)
        .SerializeToString())

    # proto stream
    chunks = []
    header = M["SerializationHeader"]()
    header.format_version = 2
    chunks.append(header.SerializeToString())
    graph = M["SerializedData"]()
    graph.pose_graph.payload = bytes(range(200))
    graph.pose_graph.number = -5
    chunks.append(graph.SerializeToString())
    hist = np.linspace(0.0, 1.0, 120).astype(np.float32)
    lx, la = prob_lo.export()
    s0 = M["SerializedData"]()
    s0.submap.submap_id.trajectory_id = 0
    s0.submap.submap_id.submap_index = 3
    s0.submap.submap_3d.local_pose.translation.x = 1.5
    s0.submap.submap_3d.local_pose.translation.z = -0.25
    s0.submap.submap_3d.local_pose.rotation.w = float(np.cos(0.3))
    s0.submap.submap_3d.local_pose.rotation.z = float(np.sin(0.3))
    s0.submap.submap_3d.num_range_data = 160
    s0.submap.submap_3d.finished = True
    fill_grid(s0.submap.submap_3d.high_resolution_hybrid_grid, px, pa)
    fill_grid(s0.submap.submap_3d.low_resolution_hybrid_grid, lx, la, resolution=0.45)
    s0.submap.submap_3d.rotational_scan_matcher_histogram.extend(float(v) for v in hist)
    chunks.append(s0.SerializeToString())
    exp["low_prob_xyz"], exp["low_prob_values"] = lx, la
    node = M["SerializedData"]()
    node.node.payload = b"node"
    chunks.append(node.SerializeToString())
    ltx, lta, ltb = tsdf_lo.export()
    s1 = M["SerializedData"]()
    s1.submap.submap_id.trajectory_id = 1
    s1.submap.submap_id.submap_index = 0
    s1.submap.submap_3d.local_pose.rotation.w = 1.0
    s1.submap.submap_3d.num_range_data = 7
    fill_grid(s1.submap.submap_3d.high_resolution_hybrid_grid_tsdf, tx, ta, tb, 0.1, 2.5, 1000.0)
    fill_grid(s1.submap.submap_3d.low_resolution_hybrid_grid_tsdf, ltx, lta, ltb, 0.45, 2.5, 1000.0)
    chunks.append(s1.SerializeToString())
    exp["low_tsdf_xyz"], exp["low_tsdf_tsd"], exp["low_tsdf_weight"] = ltx, lta, ltb
    exp["histogram"] = hist
    with open(os.path.join(HERE, "wire_map.pbstream"), "wb") as f:
        f.write(struct.pack("<Q", MAGIC))
        for c in chunks:
            z = gzip.compress(c, compresslevel=1)  # FastGzipString: gzip, best speed
            f.write(struct.pack("<Q", len(z)))
            f.write(z)
    np.savez_compressed(os.path.join(HERE, "wire_expected.npz"), **exp)
    for name in ("wire_hybrid_grid.bin", "wire_unpacked.bin", "wire_hybrid_grid_tsdf.bin", "wire_map.pbstream",
                 "wire_expected.npz"):
        print("%-28s %7d bytes" % (name, os.path.getsize(os.path.join(HERE, name))))


if __name__ == "__main__":
    main()
