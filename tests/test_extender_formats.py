"""Host-only formats around the path: topo-matrix parser, node-annotation codec, pod annotations, Allocate
reply.  Fixtures: tests/golden/*.topo.txt (F1 is the transcript of imgs/gpu_topology_on_machine.png),
annotation examples of design.md:80, 229-231, 244-245.  CPU only."""
import json
import os

import numpy as np
import pytest

import oracle_lib as ol

GOLD = os.path.join(os.path.dirname(__file__), "golden")


@pytest.mark.parametrize("name,cls", [("f1_ring", "F1_RING"), ("dgx1", "DGX1"), ("f2_pcie", "F2_PCIE"),
                                      ("f3_tree4", "F3_TREE4")])
def test_parser_reads_fixtures(gt, oracle, name, cls):
    lv, n = gt.parse_topo_matrix(open(os.path.join(GOLD, name + ".topo.txt")).read())
    want, wn, _, _ = oracle.class_topology(ol.CLS[cls])
    assert n == wn and (lv == want).all()


def test_parser_details(gt):
    # ANSI underline around the header, NIC columns, NV18 (NVSwitch) and NODE
    text = ("\x1b[4m\tGPU0\tGPU1\tGPU2\tNIC0\tCPU Affinity\tNUMA Affinity\x1b[0m\n"
            "GPU0\t X \tNV18\tNODE\tPIX\t0-31\t0\n"
            "GPU1\tNV18\t X \tSYS\tSYS\t0-31\t0\n"
            "GPU2\tNODE\tSYS\t X \tPHB\t32-63\t1\n"
            "NIC0\tPIX\tSYS\tPHB\t X \n\nLegend:\n  X = Self\n")
    lv, n = gt.parse_topo_matrix(text)
    m = lv.reshape(16, 16)
    assert n == 3
    assert m[0, 1] == gt.LV["NVS"] and m[0, 2] == gt.LV["NODE"] and m[1, 2] == gt.LV["SYS"] and m[2, 1] == gt.LV["SYS"]
    for bad in ("", "hello", "GPU0 GPU1\nGPU0 X FOO\nGPU1 FOO X\n", "GPU0 GPU1\nGPU0 X NV1\n",
                "GPU0 GPU1\nGPU0 NV1 NV1\nGPU1 NV1 X\n"):
        with pytest.raises(gt.GtError) as e:
            gt.parse_topo_matrix(bad)
        assert e.value.code == -6
    wide = " ".join("GPU%d" % i for i in range(17)) + "\n"
    with pytest.raises(gt.GtError):
        gt.parse_topo_matrix(wide)


def test_p2plink_enum_matrix_equals_parsed_text(gt, oracle):
    """The reference keeps `gpuTopology map[uint]map[uint]nvml.P2PLinkType` (design.md:57-70): the same machine read as
    that numeric enum and as `nvidia-smi topo -m` text gives the same level codes."""
    enum = {"SYS": 1, "NODE": 2, "PHB": 3, "PXB": 4, "PIX": 5, "PSB": 6, "NV1": 7, "NV2": 8, "NV3": 9, "NV4": 10}
    for name, want in enum.items():          # the rows of the table of design.md:33-44, in order
        assert gt._lib().gt_level_from_p2plink(want) == gt.LV[name]
    assert gt._lib().gt_level_from_p2plink(0) == gt.LV["SYS"]       # unknown: assume the farthest
    assert gt._lib().gt_level_from_p2plink(11) == gt.LV["NVS"] and gt._lib().gt_level_from_p2plink(12) == gt.LV["NVS"]
    assert gt._lib().gt_level_from_p2plink(13) == -1 and gt._lib().gt_level_from_p2plink(-1) == -1
    for fx in ("f1_ring", "dgx1", "f2_pcie", "f3_tree4"):
        text = open(os.path.join(GOLD, fx + ".topo.txt")).read()
        lv, n = gt.parse_topo_matrix(text)
        rows = [l.split() for l in text.splitlines() if l.split() and l.split()[0].startswith("GPU") and len(l.split()) > n]
        rows = [r for r in rows if r[1] != "GPU1"][:n]
        types = np.array([[0 if t == "X" else enum[t] for t in r[1:1 + n]] for r in rows], np.int32)
        assert (gt.levels_from_p2plink(types) == lv).all()
    with pytest.raises(gt.GtError):
        gt.levels_from_p2plink(np.full((2, 2), 99, np.int32))
    assert (gt.levels_from_p2plink(np.array([[99]], np.int32)) == 0).all()      # the diagonal is never read


def test_node_annotation_codec(gt, oracle):
    # design.md:80: "GPU_SYS_0_1: Cross CPU socket"
    lv = np.zeros(256, np.uint8)
    lv[1] = lv[16] = gt.LV["SYS"]
    assert gt.annotation_encode(lv, 2) == "GPU_SYS_0_1: Cross CPU socket\n"
    got, n = gt.annotation_decode("GPU_SYS_0_1: Cross CPU socket\nkubernetes.io/hostname: n1\n")
    assert n == 2 and (got == lv).all()
    for cls in range(6):
        clv, cn, _, _ = oracle.class_topology(cls)
        text = gt.annotation_encode(clv, cn)
        assert text.count("\n") == cn * (cn - 1) // 2
        back, bn = gt.annotation_decode(text)
        assert bn == cn and (back == clv).all()
    for bad in ("GPU_FOO_0_1: x\n", "GPU_SYS_0: x\n", "GPU_SYS_0_0: x\n", "GPU_SYS_0_2: Cross CPU socket\n"):
        with pytest.raises(gt.GtError):
            gt.annotation_decode(bad)


def test_pod_annotations_and_allocate(gt):
    # design.md:229-231
    ann = gt.bind_annotations(0b1111, 1561717704)
    assert ann == {"ALIYUN_COM_GPU_ASSIGNED": "false", "ALIYUN_COM_GPU_ASSUME_TIME": "1561717704",
                   "ALIYUN_COM_GPU_GROUP": "0,1,2,3"}
    assert gt.group_encode(0b1010_0000_0000_0001) == "0,13,15" and gt.group_decode("0,13,15") == 0b1010_0000_0000_0001
    for bad in ("0,16", "a", "1,1"):
        with pytest.raises(gt.GtError):
            gt.group_decode(bad)
    # design.md:239, 244-245
    reply, upd = gt.DevicePlugin.allocate(ann, 1561718702)
    assert reply == {"Envs": {"NVIDIA_VISIBLE_DEVICES": "0,1,2,3"}}
    assert upd == {"ALIYUN_COM_GPU_ASSIGNED": "true", "ALIYUN_COM_GPU_ASSUME_TIME": "1561718702"}
    with pytest.raises(gt.GtError):
        gt.DevicePlugin.allocate({"ALIYUN_COM_GPU_ASSIGNED": "false"}, 0)
    assert gt.DevicePlugin.resource_name == "aliyun.com/gpu"


def test_extender_policy_fixture(gt):
    want = json.load(open(os.path.join(GOLD, "extender_policy.json")))
    assert gt.EXTENDER_POLICY == want        # design.md:92-113
