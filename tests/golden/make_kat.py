"""Writes tests/golden/kat.json: the hand-derived known-answer vectors of SURVEY.md 8(c).

The expected values are LITERALS transcribed from the survey's table (derived by hand
from pkg/scheduler/score.go), not outputs of the oracle or of the GPU path.  The
reference has no test for pkg/scheduler and cannot be run here (no Go toolchain), so
these vectors are the only pin the oracle has ("parity unpinned").  Float bit patterns were worked out with
IEEE binary32 arithmetic on the literal integers named in each `cite` (e.g. float32(20)/float32(19)+float32(6)).

node = 8 device slots, Count=10, Totalmem=32768, Totalcore=100, Numa=0,
Type "NVIDIA-Tesla V100-PCIE-32GB", idle unless stated.
request = (Nums, Memreq, MemPercentagereq, Coresreq); pick = (slot, Usedmem, Usedcores).
"""
import json
import os

V100 = "NVIDIA-Tesla V100-PCIE-32GB"
A100 = "NVIDIA-NVIDIA A100-SXM4-40GB"


def node(**over):
    devs = []
    for s in range(8):
        d = dict(Used=0, Count=10, Usedmem=0, Totalmem=32768, Usedcores=0, Totalcore=100, Numa=0, Type=V100)
        for k, v in over.items():
            d[k] = v[s] if isinstance(v, (list, tuple)) else v
        devs.append(d)
    return devs


def kat(name, devices, containers, annos, fit, bits=None, picks=None, n_nodes=1, best_row=None, cite=""):
    return dict(name=name, n_nodes=n_nodes, devices=devices, containers=containers, annos=annos, fit=fit,
                score_bits=bits, picks=picks, best_row=(best_row if best_row is not None else (0 if fit else -1)),
                cite=cite)


U = 101
M370, M370X4, M290 = "MLU370-X8", "MLU370-X4", "MLU290"
mixed_types = [V100] * 4 + [A100] * 4
KATS = [
    kat("1_baseline_cfg1_64_idle_nodes", node(), [[(1, 3000, U, 30)]], {}, True, "0x41000000", [[(7, 3000, 30)]],
        n_nodes=64, best_row=63, cite="examples/nvidia/default_use.yaml; all 64 nodes tie -> highest row"),
    kat("2_two_devices_spread", node(Used=[3, 0, 5, 0, 1, 2, 9, 10]), [[(2, 3000, U, 30)]], {}, True, "0x40e00000",
        [[(3, 3000, 30), (1, 3000, 30)]], cite="equal free slots -> later slot first (stable sort + reverse walk)"),
    kat("3_three_devices", node(Used=[3, 0, 5, 0, 1, 2, 9, 10]), [[(3, 3000, U, 30)]], {}, True, "0x40c11a7c",
        [[(3, 3000, 30), (1, 3000, 30), (4, 3000, 30)]], cite="30/29 + 5"),
    kat("4_two_containers_order_persists", node(), [[(1, 1000, U, 0)], [(1, 1000, U, 0)]], {}, True, "0x41800000",
        [[(7, 1000, 0)], [(6, 1000, 0)]], cite="score.go:165 sorts in place; slot 7 moves to the front after ctr 0"),
    kat("5_exclusive", node(Used=[1, 1, 1, 0, 1, 1, 1, 1], Usedcores=[10, 10, 10, 0, 10, 10, 10, 10]),
        [[(1, 0, 100, 100)]], {}, True, "0x41000000", [[(3, 32768, 100)]], cite="score.go:128"),
    kat("6_mem_percentage_hetero", node(Totalmem=[16384] * 4 + [40960] * 2 + [81920] * 2, Used=[0, 0, 0, 0, 0, 0, 2, 2]),
        [[(2, 0, 50, 0)]], {}, True, "0x40e00000", [[(5, 20480, 0), (4, 20480, 0)]], cite="score.go:119 per device"),
    kat("7_cores0_on_full_gpu", node(Usedcores=100, Used=1), [[(1, 1000, U, 0)]], {}, False, cite="score.go:132"),
    kat("8_numa_bind", node(Numa=[0, 0, 0, 0, 1, 1, 1, 1], Used=[0, 0, 0, 0, 10, 10, 10, 0]), [[(2, 1000, U, 0)]],
        {"nvidia.com/numa-bind": "true"}, True, "0x40e00000", [[(3, 1000, 0), (2, 1000, 0)]],
        cite="score.go:99-104: slot 7 taken, then discarded at the numa change"),
    kat("8b_no_numa_bind", node(Numa=[0, 0, 0, 0, 1, 1, 1, 1], Used=[0, 0, 0, 0, 10, 10, 10, 0]), [[(2, 1000, U, 0)]],
        {}, True, "0x40e00000", [[(7, 1000, 0), (3, 1000, 0)]]),
    kat("9_use_gputype", node(Type=mixed_types), [[(1, 1000, U, 0)]], {"nvidia.com/use-gputype": "a100"}, True,
        "0x41000000", [[(7, 1000, 0)]], cite="nvidia/device.go:63-77 case-insensitive"),
    kat("10_nouse_gputype", node(Type=mixed_types), [[(1, 1000, U, 0)]], {"nvidia.com/nouse-gputype": "A100,T4"}, True,
        "0x41000000", [[(3, 1000, 0)]], cite="nvidia/device.go:78-92"),
    kat("11_more_than_devices", node(), [[(9, 1000, U, 0)]], {}, False, cite="score.go:162"),
    kat("12_no_memory_left", node(Usedmem=32000, Used=1), [[(1, 1000, U, 0)]], {}, False, cite="score.go:121"),
    kat("13_container_without_gpu", node(), [[], [(2, 2000, U, 0)]], {}, True, "0x40e00000",
        [[], [(7, 2000, 0), (6, 2000, 0)]], cite="score.go:195-198"),
    kat("14_cores_over_100", node(), [[(1, 1000, U, 101)]], {}, False, cite="score.go:110-113"),
    kat("15_all_eight", node(Used=[0, 1, 2, 3, 4, 5, 6, 7]), [[(8, 100, U, 5)]], {}, True, "0x3fc4ec4f",
        [[(s, 100, 5) for s in range(8)]], cite="80/52 + 0"),
    # Cambricon vendor: request = (Nums, Memreq, 0, 0, "MLU") -- cambricon/device.go:106-137 leaves
    # MemPercentagereq and Coresreq at Go's zero value
    kat("16_mlu370_share_needs_unused_card", node(Type=M370, Used=[1, 0, 2, 0, 0, 3, 0, 1]), [[(2, 0, 0, 0, "MLU")]], {},
        True, "0x40e00000", [[(6, 0, 0), (4, 0, 0)]], cite="cambricon/device.go:97-99: Memreq == 0 on a 370 in use fails"),
    kat("17_mlu370_share_sees_own_pick", node(Type=M370, Used=[1, 1, 1, 0, 1, 1, 1, 1]),
        [[(1, 0, 0, 0, "MLU")], [(1, 0, 0, 0, "MLU")]], {}, False,
        cite="container 0 takes slot 3; for container 1 its Used is 1 (score.go:171), so :97-99 rules it out too"),
    kat("18_mlu370_with_memory_shares_a_card", node(Type=M370, Used=[1, 1, 1, 0, 1, 1, 1, 1]),
        [[(1, 1000, 0, 0, "MLU")], [(1, 1000, 0, 0, "MLU")]], {}, True, "0x4180e38e",
        [[(3, 1000, 0)], [(3, 1000, 0)]], cite="8 + (10/9 + 7): Memreq != 0 skips :97-99"),
    kat("19_mlu290_cannot_split_memory", node(Type=M290), [[(1, 1000, 0, 0, "MLU")]], {}, False,
        cite="cambricon/device.go:94-96: Memreq != 0 on a card that is not a 370"),
    kat("20_mlu290_whole_card_in_use_still_passes", node(Type=M290, Used=1), [[(1, 0, 0, 0, "MLU")]], {}, True,
        "0x4101c71c", [[(7, 0, 0)]], cite="10/9 + 7: the Used > 0 rule is for 370 only"),
    kat("21_use_mlutype", node(Type=[M370X4] * 4 + [M370] * 4), [[(1, 0, 0, 0, "MLU")]],
        {"cambricon.com/use-mlutype": "x4"}, True, "0x41000000", [[(3, 0, 0)]], cite="cambricon/device.go:60-75"),
    kat("22_nvidia_request_on_mlu_node", node(Type=M370), [[(1, 1000, U, 0)]], {}, False, cite="score.go:73"),
    # ---- request shapes the first 23 vectors never used (derived by hand from score.go, like the rest) ----
    kat("23_memreq_wins_over_percentage", node(Used=1, Usedmem=20000), [[(1, 3000, 50, 0)]], {}, True, "0x4101c71c",
        [[(7, 3000, 0)]],
        cite="score.go:114-120: Memreq > 0 sets memreq = 3000; the percentage branch needs Memreq == 0, so 50% "
             "(16384 > 12768 free) is never computed. 10/9 + 7"),
    kat("24_negative_memreq_with_percentage", node(Used=1, Usedmem=32768), [[(1, -5, 50, 0)]], {}, True, "0x4101c71c",
        [[(7, 0, 0)]],
        cite="score.go:114-120: Memreq > 0 false and Memreq == 0 false -> memreq stays 0; 0 free >= 0 fits, Usedmem 0"),
    kat("25_negative_coresreq_then_zero_cores", node(Used=1, Usedcores=100), [[(1, 1000, U, -10)], [(1, 1000, U, 0)]],
        {}, True, "0x4182e38e", [[(7, 1000, -10)], [(7, 1000, 0)]],
        cite="score.go:124: 0 free cores >= -10; :172 Usedcores += -10 -> 90. Container 1 (cores 0): slots 0-6 are "
             "full (:132), slot 7 (now first in the sort, 8 free) is reached last and fits: (10/9 + 7) + (10/8 + 7)"),
    kat("26_used_exceeds_count", node(Used=[12, 11, 10, 9, 12, 11, 10, 12]), [[(1, 1000, U, 0)]], {}, True, "0x41880000",
        [[(3, 1000, 0)]], cite="score.go:107 Count <= Used skips seven devices; 10/1 + 7"),
    kat("27_totalcore_zero_cores_zero", node(Totalcore=0, Used=1), [[(1, 1000, U, 0)]], {}, True, "0x4101c71c",
        [[(7, 1000, 0)]], cite="score.go:124 0-0 < 0 false; :128 needs Totalcore == 100; :132 needs Totalcore != 0"),
    kat("27b_totalcore_zero_cores_100", node(Totalcore=0), [[(1, 1000, U, 100)]], {}, False,
        cite="score.go:124: 0 - 0 < 100 on every device"),
    kat("28_numa_bind_type_mismatch_inside_group",
        node(Numa=[0, 0, 0, 0, 1, 1, 1, 1], Type=[V100] * 5 + [A100] + [V100] * 2), [[(3, 1000, U, 0)]],
        {"nvidia.com/numa-bind": "true", "nvidia.com/use-gputype": "v100"}, True, "0x40c00000",
        [[(7, 1000, 0), (6, 1000, 0), (4, 1000, 0)]],
        cite="score.go:94-104: slot 5 fails the type check and is skipped BEFORE the numa test, so the numa-1 group "
             "is not broken: 7, 6, 4. 30/30 + 5"),
    kat("29_second_request_sees_first_take", node(Used=[1] * 7 + [0], Usedmem=[32768] * 7 + [0]),
        [[(1, 1000, U, 0, "NVIDIA"), (1, 1000, U, 100, "NVIDIA")]], {}, False,
        cite="score.go:168-174: request 0 takes slot 7 (the only one with memory) -> Used 1; request 1 wants a card "
             "exclusively (:128 Used > 0) -> nothing fits, the container fails"),
    kat("29b_second_request_shares_the_card", node(Used=[1] * 7 + [0], Usedmem=[32768] * 7 + [0]),
        [[(1, 1000, U, 0, "NVIDIA"), (1, 1000, U, 50, "NVIDIA")]], {}, True, "0x40e1af28",
        [[(7, 1000, 0), (7, 1000, 50)]],
        cite="same, but 50 cores: slot 7 again. total 10 + 10, free 10 + 9 (:170 after the first Used++), sums 2: 20/19 + 6"),
    kat("30_two_vendors_in_one_container", node(Type=[V100] * 4 + [M370] * 4),
        [[(1, 0, 0, 0, "MLU"), (1, 1000, U, 0, "NVIDIA")]], {}, True, "0x40e00000", [[(7, 0, 0), (3, 1000, 0)]],
        cite="MLU request takes slot 7 (Used 1 -> sorts first); the NVIDIA request walks 6, 5, 4 (type mismatch, "
             "score.go:73) and takes slot 3. 20/20 + 6"),
    kat("31_three_device_node", [dict(Used=u, Count=4, Usedmem=0, Totalmem=16000, Usedcores=0, Totalcore=100, Numa=0, Type=V100)
                                 for u in (1, 3, 0)], [[(2, 1000, U, 0)]], {}, True, "0x40092492", [[(2, 1000, 0), (0, 1000, 0)]],
        cite="len(Devices) = 3. Count-Used = 3, 1, 4 -> ascending order slots 1, 0, 2 (score.go:45-50); the walk from the "
             "end (:92) takes slot 2 (total 4, free 4) and slot 0 (total 8, free 7): 8/7 + (3 - 2)"),
    kat("32_second_container_finds_nothing", node(Count=[10] * 7 + [1], Used=[10] * 7 + [0]),
        [[(1, 1000, U, 0)], [(1, 1000, U, 0)]], {}, False,
        cite="slots 0-6 have Count <= Used (score.go:107); container 0 takes the one share of slot 7 (:168-174), container "
             "1 finds no device (:175-177) -> the node is no candidate (:200-206)"),
    kat("33_percentage_truncates", node(Totalmem=1001, Usedmem=[672, 672, 671, 672, 672, 672, 672, 672]),
        [[(1, 0, 33, 0)]], {}, True, "0x41000000", [[(2, 330, 0)]],
        cite="score.go:119: 1001 * 33 / 100 = 330 in integers; free memory is 329 everywhere but on slot 2 (330): the walk "
             "passes slots 7..3 (:121) and takes slot 2. 10/10 + 7"),
]

# string-policy truth table (nvidia/device.go:62-105), each line derived by hand
POLICY = [
    # (annos, card type, expected pass)
    ({}, V100, True),
    ({"nvidia.com/use-gputype": "v100"}, V100, True),
    ({"nvidia.com/use-gputype": "A100"}, V100, False),
    ({"nvidia.com/use-gputype": "A100,v100"}, V100, True),
    ({"nvidia.com/use-gputype": ""}, V100, True),            # strings.Contains(x, "") is true
    ({"nvidia.com/use-gputype": "A100,"}, V100, True),       # trailing comma -> empty token matches
    ({"nvidia.com/nouse-gputype": "A100,"}, V100, False),    # ... and excludes everything for nouse
    ({"nvidia.com/nouse-gputype": "a100"}, A100, False),
    ({"nvidia.com/nouse-gputype": "a100"}, V100, True),
    ({"nvidia.com/use-gputype": "v100", "nvidia.com/nouse-gputype": "v100"}, V100, True),   # use wins
    ({"nvidia.com/use-gputype": "a100", "nvidia.com/nouse-gputype": "t4"}, V100, False),    # nouse not consulted
    ({"nvidia.com/use-gputype": "Tesla V100"}, V100, True),
    ({"nvidia.com/use-gputype": "tesla-v100"}, V100, False),
]
NUMA = [("true", 1), ("True", 1), ("TRUE", 1), ("t", 1), ("T", 1), ("1", 1), ("false", 0), ("0", 0), ("yes", 0),
        ("tRuE", 0), ("", 0), (None, 0)]

# pkg/util/util_test.go:42-56 (TestPodDevicesCoding) -- the one golden string the reference's
# own tests imply for the result wire format, via util.go:110-126
ENCODE = [
    ([[("UUID1", "Type1", 1000, 30)], [], [("UUID1", "Type1", 1000, 30)]], "UUID1,Type1,1000,30:;;UUID1,Type1,1000,30:"),
    ([], ""),
    ([[]], ""),
]

if __name__ == "__main__":
    out = os.path.join(os.path.dirname(os.path.abspath(__file__)), "kat.json")
    with open(out, "w") as f:
        json.dump(dict(kats=KATS, policy=POLICY, numa=NUMA, encode=ENCODE), f, indent=1)
    print("wrote", out, len(KATS), "KATs")
