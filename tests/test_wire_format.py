"""Grid wire format: the library's decoder / encoder for proto::HybridGrid,
proto::HybridGridTSDF and the .pbstream container against fixtures ENCODED BY GOOGLE'S PROTOBUF
RUNTIME (tests/golden/make_wire_golden.py). CPU tests cover the host-side decode; GPU tests the
upload (HybridGrid(proto) / HybridGridTSDF(proto) semantics vs the oracle) and ToProto."""
import os

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
G = os.path.join(HERE, "golden")


def read(name):
    return open(os.path.join(G, name), "rb").read()


@pytest.fixture(scope="module")
def expected():
    return np.load(os.path.join(G, "wire_expected.npz"))


@pytest.fixture(scope="module")
def hg():
    import hectorgrapher_b200 as hg
    hg.lib()
    return hg


def test_decode_hybrid_grid_packed_and_unpacked(hg, expected):
    for name in ("wire_hybrid_grid.bin", "wire_unpacked.bin"):
        info, xyz, a, b = hg.grid_proto_decode(read(name), tsdf=False)
        assert info["resolution"] == np.float32(0.1) and info["num_voxels"] == len(expected["prob_xyz"])
        assert (xyz == expected["prob_xyz"]).all() and (a == expected["prob_values"]).all() and b is None
    assert (xyz[-1] == [-8000, 8000, -1]).all()


def test_decode_hybrid_grid_tsdf(hg, expected):
    info, xyz, a, b = hg.grid_proto_decode(read("wire_hybrid_grid_tsdf.bin"), tsdf=True)
    assert info["resolution"] == np.float32(0.1) and info["max_weight"] == 1000.0
    assert info["relative_truncation_distance"] == np.float32(2.5) * np.float32(0.1)  # ToProto's quirk
    assert (xyz == expected["tsdf_xyz"]).all()
    assert (a == expected["tsdf_tsd"]).all() and (b == expected["tsdf_weight"]).all()
    assert (a >= 0x8000).all()  # stored TSD values carry the update marker through the wire format


def test_decode_rejects_malformed_messages(hg):
    good = read("wire_hybrid_grid.bin")
    with pytest.raises(hg.HgError):
        hg.grid_proto_decode(good[:len(good) // 2], tsdf=False)      # truncated in a packed field
    with pytest.raises(hg.HgError):
        hg.grid_proto_decode(b"\x0d\xcd\xcc\xcc\x3d\x1a\x02\x02\x04", tsdf=False)  # 2 x indices, no values
    with pytest.raises(hg.HgError):
        hg.grid_proto_decode(b"\x1a\x00", tsdf=False)                 # no resolution
    info, xyz, a, _ = hg.grid_proto_decode(b"\x0d\xcd\xcc\xcc\x3d", tsdf=False)  # an empty grid is fine
    assert info["num_voxels"] == 0


def test_pbstream_walk_without_upload(hg, expected):
    submaps = [(d, h) for d, _, _, h in hg.PbStream(os.path.join(G, "wire_map.pbstream"), upload=False)]
    assert len(submaps) == 2  # the header, the pose-graph stand-in and the node are skipped
    d0, h0 = submaps[0]
    assert (d0["trajectory_id"], d0["submap_index"], d0["num_range_data"], d0["finished"]) == (0, 3, 160, True)
    assert np.allclose(d0["local_pose"], [1.5, 0, -0.25, np.cos(0.3), 0, 0, np.sin(0.3)], atol=0)
    assert (d0["high_is_tsdf"], d0["low_is_tsdf"]) == (0, 0)
    assert (h0 == expected["histogram"]).all()
    d1, h1 = submaps[1]
    assert (d1["trajectory_id"], d1["submap_index"], d1["num_range_data"], d1["finished"]) == (1, 0, 7, False)
    assert (d1["local_pose"] == [0, 0, 0, 1, 0, 0, 0]).all()
    assert (d1["high_is_tsdf"], d1["low_is_tsdf"]) == (1, 1) and len(h1) == 0
    with pytest.raises(hg.HgError):
        hg.PbStream(os.path.join(G, "wire_hybrid_grid.bin"))  # not a proto stream


def sorted_voxels(xyz, a, b=None):
    order = np.lexsort((xyz[:, 0], xyz[:, 1], xyz[:, 2]))
    return xyz[order], a[order], (None if b is None else b[order])


def oracle_from_proto(oracle, info, xyz, a, b, tsdf):
    """The proto constructors on the oracle: SetProbability(ValueToProbability(v)) /
    SetCell(ValueToTSD(t), ValueToWeight(w)) — hybrid_grid.h:67-71, hybrid_grid_tsdf.h:78-82."""
    if not tsdf:
        g = oracle.Grid.prob(info["resolution"])
        table = oracle.value_table(0.1, 0.1, 0.9)
        g.set_probability(xyz, table[a])
        return g
    g = oracle.Grid.tsdf(info["resolution"], info["relative_truncation_distance"], info["max_weight"])
    max_tsd = np.float32(info["relative_truncation_distance"]) * np.float32(info["resolution"])
    g.set_cell(xyz, oracle.value_table(-max_tsd, -max_tsd, max_tsd)[a],
               oracle.value_table(0.0, 0.0, info["max_weight"])[b])
    return g


@pytest.mark.gpu
@pytest.mark.parametrize("name,tsdf", [("wire_hybrid_grid.bin", False), ("wire_unpacked.bin", False),
                                       ("wire_hybrid_grid_tsdf.bin", True)])
def test_from_proto_matches_the_reference_constructors(oracle, hg, name, tsdf):
    data = read(name)
    g = hg.Grid.from_proto(data, tsdf)
    info, xyz, a, b = hg.grid_proto_decode(data, tsdf)
    o = oracle_from_proto(oracle, info, xyz, a, b, tsdf)
    got, want = sorted_voxels(*g.export()), sorted_voxels(*o.export())
    assert (got[0] == want[0]).all() and (got[1] == want[1]).all()
    if tsdf:
        assert (got[2] == want[2]).all()
    assert g.grid_size == o.grid_size
    what = "tsd" if tsdf else "probability"
    assert (g.lookup(what, xyz[::3]) == o.lookup(what, xyz[::3])).all()


@pytest.mark.gpu
def test_to_proto_round_trip(hg, expected):
    """ToProto -> bytes -> decode gives the grid's voxels; for a probability grid a second
    from_proto reproduces the grid; a TSDF grid comes back with max_tsd * resolution as its
    truncation distance, exactly like the reference's ToProto / constructor pair."""
    g = hg.Grid.from_proto(read("wire_hybrid_grid.bin"), False)
    data = g.to_proto()
    info, xyz, a, _ = hg.grid_proto_decode(data, False)
    e = sorted_voxels(*g.export()[:2])
    d = sorted_voxels(xyz, a)
    assert (d[0] == e[0]).all() and (d[1] == e[1]).all() and info["resolution"] == np.float32(0.1)
    g2 = hg.Grid.from_proto(data, False)
    e2 = sorted_voxels(*g2.export()[:2])
    assert (e2[0] == e[0]).all() and (e2[1] == e[1]).all()
    # python-protobuf parses what the library wrote
    import sys
    sys.path.insert(0, G)
    import make_wire_golden as mw
    M = mw.build_messages()
    msg = M["HybridGrid"]()
    msg.ParseFromString(data)
    assert len(msg.values) == len(a) and abs(msg.resolution - 0.1) < 1e-7
    assert sorted(zip(msg.z_indices, msg.y_indices, msg.x_indices, msg.values)) == \
        sorted(zip(xyz[:, 2].tolist(), xyz[:, 1].tolist(), xyz[:, 0].tolist(), a.tolist()))

    t = hg.Grid.tsdf(0.1, 2.5, 1000.0, expected["tsdf_xyz"], expected["tsdf_tsd"], expected["tsdf_weight"])
    tdata = t.to_proto()
    tinfo, txyz, ta, tb = hg.grid_proto_decode(tdata, True)
    assert tinfo["relative_truncation_distance"] == np.float32(2.5) * np.float32(0.1)
    assert tinfo["max_weight"] == 1000.0 and len(ta) == len(expected["tsdf_tsd"])
    tmsg = M["HybridGridTSDF"]()
    tmsg.ParseFromString(tdata)
    assert len(tmsg.values_weight) == len(tb)
    # byte-identical to what protobuf itself emits for the same content (same voxel order)
    again = mw.fill_grid(M["HybridGridTSDF"](), txyz, ta, tb, 0.1, float(tinfo["relative_truncation_distance"]),
                         1000.0).SerializeToString()
    assert again == tdata


@pytest.mark.gpu
def test_pbstream_submaps_feed_the_matcher(oracle, hg, expected):
    submaps = list(hg.PbStream(os.path.join(G, "wire_map.pbstream")))
    (d0, hi0, lo0, h0), (d1, hi1, lo1, h1) = submaps
    assert not hi0.is_tsdf and not lo0.is_tsdf and hi1.is_tsdf and lo1.is_tsdf
    e = sorted_voxels(*lo0.export()[:2])
    w = sorted_voxels(expected["low_prob_xyz"], expected["low_prob_values"])
    assert (e[0] == w[0]).all() and (e[1] == w[1]).all()
    e = sorted_voxels(*lo1.export())
    w = sorted_voxels(expected["low_tsdf_xyz"], expected["low_tsdf_tsd"], expected["low_tsdf_weight"])
    assert (e[0] == w[0]).all() and (e[1] == w[1]).all() and (e[2] == w[2]).all()
    # a pyramid and a matcher can be built straight from the stream's grids
    stack = hg.Stack(hi1, 4, 3)
    assert len(stack.export(0)[0]) > 0
    hg.Fcsm(hi1, lo1, np.zeros(8, np.float32), branch_and_bound_depth=4, full_resolution_depth=3)


@pytest.mark.parametrize("seed", range(6))
def test_decode_random_messages_against_protobuf(hg, seed):
    """Random grids (the whole +-8192 index range, value 0 and marker-carrying values, empty grids,
    packed and unpacked encodings, an unknown field in front) encoded by the protobuf runtime and
    decoded by the library; CPU only."""
    import sys
    sys.path.insert(0, G)
    import make_wire_golden as mw
    M = mw.build_messages()
    rng = np.random.RandomState(100 + seed)
    n = [0, 1, 7, 300, 2000, 5000][seed]
    xyz = rng.randint(-8192, 8192, (n, 3)).astype(np.int32)
    a = rng.randint(0, 65536, n).astype(np.uint16)
    b = rng.randint(0, 65536, n).astype(np.uint16)
    res = float(np.float32(rng.uniform(0.01, 2.0)))
    for name in ("HybridGrid", "HybridGridUnpacked"):
        data = mw.fill_grid(M[name](), xyz, a, resolution=res).SerializeToString()
        data = b"\x92\x03\x03abc" + data  # field 50, length-delimited: skipped
        info, gx, ga, gb = hg.grid_proto_decode(data, tsdf=False)
        assert info["num_voxels"] == n and info["resolution"] == np.float32(res)
        assert (gx == xyz).all() and (ga == a).all()
    data = mw.fill_grid(M["HybridGridTSDF"](), xyz, a, b, res, 0.25, 123.5).SerializeToString()
    info, gx, ga, gb = hg.grid_proto_decode(data, tsdf=True)
    assert info["relative_truncation_distance"] == 0.25 and info["max_weight"] == 123.5
    assert (gx == xyz).all() and (ga == a).all() and (gb == b).all()
    # negative int32 values are 10-byte varints on the wire; the decoder narrows them like protobuf does
    msg = M["HybridGrid"]()
    msg.resolution = 0.5
    msg.x_indices.append(1), msg.y_indices.append(-2), msg.z_indices.append(3), msg.values.append(-1)
    info, gx, ga, _ = hg.grid_proto_decode(msg.SerializeToString(), tsdf=False)
    assert (gx == [[1, -2, 3]]).all() and ga[0] == 0xFFFF
