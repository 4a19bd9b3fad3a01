"""The C-ABI library loads and exports every symbol include/vgpuscore.h declares (no GPU needed)."""
import ctypes
import os
import re

import numpy as np
import pytest

import vgpu_b200 as v
from helpers import load_kats

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "vgpuscore.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(vgs_[a-z0-9_]+)\s*\(", src)))


def test_exports(built):
    lib = v.load_library()
    names = declared_symbols()
    assert len(names) >= 20
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/vgpuscore.h but not exported"
    assert set(names) == set(v.capi.EXPORTS)
    assert lib.vgs_abi_version() == 2
    assert lib.vgs_strerror(0) == b"ok"


def test_struct_layouts(built):
    assert ctypes.sizeof(v.capi.VgsConfig) == 48
    assert ctypes.sizeof(v.capi.VgsTuning) == 72 and ctypes.sizeof(v.capi.VgsPlan) == 176
    assert ctypes.sizeof(v.capi.VgsTiming) == 32
    assert v.capi.REQ_DTYPE.itemsize == 32 and v.capi.REQ_DTYPE.fields["typemask"][1] == 24
    assert v.capi.DELTA_DTYPE.itemsize == 20


def test_init_without_gpu_fails_loudly(built):
    import torch
    if torch.cuda.is_available():
        return
    try:
        v.Context(max_nodes=16)
    except v.VgsError as e:
        assert e.code == v.capi.VGS_ECUDA
    else:
        raise AssertionError("vgs_init must fail without a GPU (no CPU fallback)")


def test_host_policy_matches_truth_table_and_oracle(built):
    import oracle
    K = load_kats()
    for annos, card, want in K["policy"]:
        m = v.capi.host_typemask("NVIDIA", annos, [card])
        assert bool(m & 1) == want, (annos, card)
        assert m == oracle.typemask(annos, "NVIDIA", [card])[0]
    for val, want in K["numa"]:
        annos = {} if val is None else {"nvidia.com/numa-bind": val}
        assert v.capi.host_numa_bind(annos) == want, val
    names = [t for t, _ in v.workload.GPU_TYPES]
    _, batch = v.workload.make_config("cfg3", 8, 512)
    for an in batch.annos:
        assert v.capi.host_typemask("NVIDIA", an, names) == oracle.typemask(an, "NVIDIA", names)[0]
    assert v.capi.host_typemask("MLU", {}, ["MLU370"]) == 0  # Cambricon has its own entry point
    mlu = ["MLU370-X8", "MLU290", "NVIDIA-x"]
    for an in ({}, {"cambricon.com/use-mlutype": "mlu290"}, {"cambricon.com/nouse-mlutype": "x8,t4"},
               {"cambricon.com/use-mlutype": "X8,", "cambricon.com/nouse-mlutype": "x8"}):
        for memreq in (0, 1000):
            m, fl = v.capi.host_typemask_mlu(memreq, an, mlu)
            om, of = oracle.typemask(an, "MLU", mlu, memreq=memreq)
            assert (m, fl == v.capi.VGS_REQ_IDLE_SHARE) == (om, of == 2), (an, memreq)
    assert [v.capi.host_type_flags(t) for t in mlu] == [v.capi.VGS_TYPE_IDLE_SHARE, 0, 0]


def test_encode_pod_devices_golden(built):
    for ctrs, want in load_kats()["encode"]:
        assert v.capi.encode_pod_devices([[tuple(d) for d in c] for c in ctrs]) == want


def test_pack_batch_layout(built):
    _, batch = v.workload.make_config("cfg4", 8, 50)
    pods, reqs = v.capi.pack_batch(batch, [t for t, _ in v.workload.GPU_TYPES])
    assert len(pods) == 50 and int(pods["n_reqs"].sum()) == len(reqs)
    for p in range(50):
        r0, n = int(pods["req_begin"][p]), int(pods["n_reqs"][p])
        ctr = reqs["ctr"][r0:r0 + n]
        assert ctr[0] == 0 and (np.diff(ctr) >= 0).all()
    numa = np.array([1 if a else 0 for a in batch.annos])
    assert numa.sum() > 0


def test_decode_node_devices(built):
    """util.DecodeNodeDevices (util.go:68-99) on strings in util.EncodeNodeDevices' format (:101-108)"""
    anno = "GPU-aa,10,32768,100,NVIDIA-Tesla V100-PCIE-32GB,0,true:GPU-bb,4,16160,100,NVIDIA-Tesla T4,1,false:"
    devs = v.capi.decode_node_devices(anno)
    assert devs == [
        dict(ID="GPU-aa", Count=10, Devmem=32768, Devcore=100, Type="NVIDIA-Tesla V100-PCIE-32GB", Numa=0, Health=True),
        dict(ID="GPU-bb", Count=4, Devmem=16160, Devcore=100, Type="NVIDIA-Tesla T4", Numa=1, Health=False)]
    # Atoi errors are ignored by the reference (-> 0); entries without ',' are skipped
    devs = v.capi.decode_node_devices("x,abc,-5,+7,T,,maybe::junk:")
    assert devs == [dict(ID="x", Count=0, Devmem=-5, Devcore=7, Type="T", Numa=0, Health=False)]
    # strconv.Atoi = ParseInt(s, 10, 0): a range error clamps to the int64 limit, which int32(x) then truncates
    # (max int64 -> -1, min int64 -> 0); the scan stops at the overflow, before it can see a later bad byte
    devs = v.capi.decode_node_devices("y,99999999999999999999,-99999999999999999999,99999999999999999999x,T,4294967297,1:")
    assert devs == [dict(ID="y", Count=-1, Devmem=0, Devcore=-1, Type="T", Numa=1, Health=True)]
    for bad in ("no separators", "a,b,c:", "a,1,2,3,T,0,true,extra:"):
        try:
            v.capi.decode_node_devices(bad)
        except v.VgsError as e:
            assert e.code == v.capi.VGS_EINVAL
        else:
            raise AssertionError(bad)
    # decoded columns are exactly what the snapshot packer consumes
    snap = v.Snapshot.from_nodes([[dict(Count=d["Count"], Totalmem=d["Devmem"], Totalcore=d["Devcore"], Type=d["Type"],
                                        Numa=d["Numa"]) for d in v.capi.decode_node_devices(anno)]])
    assert snap.planes[:, 0, 1].tolist() == [0, 4, 0, 16160, 0, 100, 1, 1]


def test_pod_devices_coding_round_trips(built):
    """pkg/util/util_test.go:26-56 (TestEmptyContainerDevicesCoding, TestEmptyPodDeviceCoding,
    TestPodDevicesCoding): Encode then Decode gives the value back"""
    enc, dec = v.capi.encode_pod_devices, v.capi.decode_pod_devices
    assert dec(enc([])) == []  # TestEmptyPodDeviceCoding (and the empty ContainerDevices of :26-32)
    pd1 = [[("UUID1", "Type1", 1000, 30)], [], [("UUID1", "Type1", 1000, 30)]]  # TestPodDevicesCoding
    assert enc(pd1) == "UUID1,Type1,1000,30:;;UUID1,Type1,1000,30:"
    assert dec(enc(pd1)) == pd1
    rng = np.random.default_rng(5)
    for _ in range(50):
        pd = [[(f"GPU-{rng.integers(1 << 30):x}", str(rng.choice(["NVIDIA", "MLU", "DCU"])), int(rng.integers(0, 1 << 20)),
                int(rng.integers(0, 101))) for _ in range(int(rng.integers(0, 5)))] for _ in range(int(rng.integers(1, 6)))]
        if any(pd):  # a pod of only empty containers encodes to ";;" pieces and still round-trips...
            assert dec(enc(pd)) == pd
    assert dec(enc([[], []])) == [[], []]
    assert dec(enc([[]])) == []  # ... except the single empty container: "" decodes to no containers (util.go:159-161)


def test_decode_pod_devices_follows_the_go_parser(built):
    dec = v.capi.decode_pod_devices
    # fewer than 4 fields: DecodeContainerDevices errors, DecodePodDevices returns an EMPTY list and nil (util.go:144-146,167-169)
    assert dec("a,b,1:") == [] and dec("u,t,1,2:;x,y:") == []
    # pieces without a comma are skipped (:141); extra fields are ignored
    assert dec("junk:u,t,5,6,extra,more:") == [[("u", "t", 5, 6)]]
    # strconv.ParseInt(s, 10, 32) with the error dropped: syntax error -> 0, range error -> the int32 limit,
    # and the scan stops at the overflow before it can see a later bad byte
    assert dec("u,t,12x,-:") == [[("u", "t", 0, 0)]]
    assert dec("u,t,2147483648,-2147483649:") == [[("u", "t", 2147483647, -2147483648)]]
    assert dec("u,t,99999999999x,2147483648x:") == [[("u", "t", 2147483647, 0)]]
    assert dec("u,t,+7,-0:") == [[("u", "t", 7, 0)]]
    with pytest.raises(v.VgsError):
        dec("a,b,1,2:" * 5, max_dev=4)


def test_header_is_plain_c99_and_links(built, tmp_path):
    """the cgo shim includes vgpuscore.h as C: the header must compile as C99 with no C++ leaking in, and a
    C program that names every declared entry point must link against the in-tree library"""
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    hdr = os.path.join(root, "include", "vgpuscore.h")
    names = sorted(set(re.findall(r"\b(vgs_[a-z0-9_]+)\s*\(", open(hdr).read())))
    assert set(names) == set(v.capi.EXPORTS)
    src = tmp_path / "abi.c"
    body = "\n".join(f"    p[{i}] = (fn){n};" for i, n in enumerate(names))
    src.write_text(f'#include "vgpuscore.h"\n#include <stdio.h>\ntypedef void (*fn)(void);\nint main(void) {{\n    fn p[{len(names)}];\n{body}\n'
                   f'    printf("%d %d %d\\n", vgs_abi_version(), (int)sizeof(vgs_req), (int)sizeof(vgs_result));\n'
                   f'    return p[0] == 0;\n}}\n')
    exe = tmp_path / "abi"
    libdir = os.path.dirname(v.capi.LIB_PATH)
    subprocess.check_call(["gcc", "-std=c99", "-pedantic", "-Wall", "-Werror", "-I", os.path.join(root, "include"),
                           str(src), "-o", str(exe), "-L", libdir, "-lvgpuscore", f"-Wl,-rpath,{libdir}"])
    out = subprocess.check_output([str(exe)]).decode().split()
    assert int(out[0]) >= 1 and out[1:] == ["32", "16"]


def test_narrow_planes_pick_the_narrowest_type():
    """the packer side of vgs_snapshot_upload_narrow: every plane in the narrowest of int8 / int16 / int32 that holds
    its values, values unchanged (the GPU sign-extends them back to the int32 planes)"""
    planes = np.zeros((8, 5, 4), np.int32)
    planes[0, 1, 2], planes[0, 0, 0] = 127, -128          # int8 to the edge
    planes[1, 1, 1] = 128                                  # one past: int16
    planes[2, 0, 3] = -32769                               # one past int16: int32
    planes[3] = 183359                                     # memory in MB
    planes[4, 2, 2] = 32767
    planes[5, 4, 3] = -129
    ndev = np.array([0, 4, 2, 4, 1], np.int32)
    arrs = v.capi.narrow_planes(v.Snapshot(planes, ndev, ["NVIDIA-A100"]))
    assert [a.dtype.itemsize for a in arrs] == [1, 2, 4, 4, 2, 2, 1, 1, 1]
    for f in range(8):
        assert np.array_equal(arrs[f].astype(np.int32), planes[f]) and arrs[f].flags["C_CONTIGUOUS"]
    assert np.array_equal(arrs[8].astype(np.int32), ndev)
    empty = v.capi.narrow_planes(v.Snapshot(np.zeros((8, 0, 4), np.int32), np.zeros(0, np.int32), ["x"]))
    assert [a.dtype.itemsize for a in empty] == [1] * 9
