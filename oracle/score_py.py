"""TEST INFRASTRUCTURE ONLY -- R0: a third restatement of the Filter scoring path, in plain Python, read from the Go
again and sharing no code with score_ref.c (R1 / R2).  It exists to triangulate the oracle: three restatements in two
languages have to agree bit for bit on small random cases (tests/test_oracle_triangle.py).  Slow by design (Python
objects, one struct per device, strings compared per device and request): small cases only.  PARITY UNPINNED like the
rest of oracle/ -- the reference cannot be run here.

Follows, line by line:
  pkg/scheduler/score.go:45-50    DeviceUsageList.Less
  pkg/scheduler/score.go:71-84    checkType
  pkg/scheduler/score.go:86-152   fitInCertainDevice
  pkg/scheduler/score.go:154-181  fitInDevices
  pkg/scheduler/score.go:183-214  calcScore
  pkg/scheduler/scheduler.go:386-387  sort.Sort(nodeScores), last element (canonical tie: highest score, highest row)
  pkg/device/nvidia/device.go:62-112, hygon/device.go:45-84, cambricon/device.go:60-104   CheckType of the vendors
  Go 1.19 sort.Sort (insertion sort for n <= 12), strings.ToUpper / strings.Map / unicode.ToUpper (Unicode 13.0.0),
  strconv.ParseBool, int32 wrap-around, float32 arithmetic.
"""
from types import SimpleNamespace

import numpy as np

MEMPCT_UNSET = 101

# ---- Go scalars ----------------------------------------------------------------------------------------------------


def i32(x: int) -> int:
    """wrap to a signed 32-bit value, as Go's int32 arithmetic does"""
    x &= 0xFFFFFFFF
    return x - 0x100000000 if x & 0x80000000 else x


def go_div(a: int, b: int) -> int:
    """Go's integer division truncates toward zero"""
    q = abs(a) // abs(b)
    return i32(q if (a < 0) == (b < 0) else -q)


F32 = np.float32

# ---- Go strings ----------------------------------------------------------------------------------------------------

_LATER = {0x2C5F, 0xA7C1, 0xA7D1, 0xA7D7, 0xA7D9} | set(range(0x10597, 0x105BD))  # case pairs newer than Unicode 13.0.0
_FFFD = "�".encode()


def _unicode_to_upper(cp: int) -> int:
    """unicode.ToUpper: simple (one-to-one) uppercase mapping"""
    if cp in _LATER:
        return cp
    c = chr(cp)
    u = c.upper()
    if len(u) == 1:
        return ord(u)
    t = c.title()  # (the characters whose full uppercase is several letters: the simple one is the title case, if single)
    return ord(t) if len(t) == 1 and t != c else cp


def strings_to_upper(s: bytes) -> bytes:
    """strings.ToUpper: ASCII byte by byte, else strings.Map(unicode.ToUpper, s) -- a range loop over the string, in
    which a byte that starts no valid encoding is (U+FFFD, width 1)"""
    if all(b < 0x80 for b in s):
        return bytes(b - 32 if 0x61 <= b <= 0x7A else b for b in s)
    out, i = bytearray(), 0
    while i < len(s):
        b0 = s[i]
        need = 1 if b0 < 0x80 else 2 if 0xC2 <= b0 <= 0xDF else 3 if 0xE0 <= b0 <= 0xEF else 4 if 0xF0 <= b0 <= 0xF4 else 0
        ch = None
        if need:
            try:
                ch = s[i:i + need].decode("utf-8")
            except UnicodeDecodeError:
                ch = None
        if ch is None or len(ch) != 1:
            out += _FFFD
            i += 1
        else:
            out += chr(_unicode_to_upper(ord(ch))).encode()
            i += need
    return bytes(out)


def parse_bool(s: str):
    """strconv.ParseBool -> (value, ok)"""
    if s in ("1", "t", "T", "TRUE", "true", "True"):
        return True, True
    if s in ("0", "f", "F", "FALSE", "false", "False"):
        return False, True
    return False, False


# ---- the vendors' CheckType ------------------------------------------------------------------------------------------


def _check_list_type(annos: dict, cardtype: str, use_key: str, nouse_key: str) -> bool:
    """checkGPUtype / checkDCUtype / checkMLUtype: the three are the same function over different annotation keys"""
    card_up = strings_to_upper(cardtype.encode())
    if use_key in annos:
        inuse = annos[use_key]
        if "," not in inuse:
            if strings_to_upper(inuse.encode()) in card_up:
                return True
        else:
            for val in inuse.split(","):
                if strings_to_upper(val.encode()) in card_up:
                    return True
        return False
    if nouse_key in annos:
        nouse = annos[nouse_key]
        if "," not in nouse:
            if strings_to_upper(nouse.encode()) in card_up:
                return False
        else:
            for val in nouse.split(","):
                if strings_to_upper(val.encode()) in card_up:
                    return False
        return True
    return True


def _assert_numa(annos: dict) -> bool:
    if "nvidia.com/numa-bind" in annos:
        enforce, ok = parse_bool(annos["nvidia.com/numa-bind"])
        if ok and enforce:
            return True
    return False


def _nvidia_check_type(annos, d, n):
    if n.Type == "NVIDIA":
        return True, _check_list_type(annos, d.Type, "nvidia.com/use-gputype", "nvidia.com/nouse-gputype"), _assert_numa(annos)
    return False, False, False


def _hygon_check_type(annos, d, n):
    if n.Type == "DCU":
        return True, _check_list_type(annos, d.Type, "hygon.com/use-dcutype", "hygon.com/nouse-dcutype"), False
    return False, False, False


def _cambricon_check_type(annos, d, n):
    if "MLU" in n.Type:
        if "370" not in d.Type and n.Memreq != 0:
            return True, False, False
        if "370" in d.Type and n.Memreq == 0 and d.Used > 0:
            return True, False, False
        return True, _check_list_type(annos, d.Type, "cambricon.com/use-mlutype", "cambricon.com/nouse-mlutype"), False
    return False, False, False


# device.GetDevices() is a map: its order is random in Go, and immaterial -- at most one vendor claims a request type
# among "NVIDIA", "DCU" and "MLU"
_VENDORS = (_cambricon_check_type, _nvidia_check_type, _hygon_check_type)


def check_type(annos, d, n):
    if n.Type not in d.Type:  # strings.Contains(d.Type, n.Type)
        return False, False
    for vendor in _VENDORS:
        found, passed, numa_assert = vendor(annos, d, n)
        if found:
            return passed, numa_assert
    return False, False


# ---- score.go ----------------------------------------------------------------------------------------------------------


class DeviceUsage:
    __slots__ = ("Id", "Used", "Count", "Usedmem", "Totalmem", "Totalcore", "Usedcores", "Numa", "Type")


def _less(a: DeviceUsage, b: DeviceUsage) -> bool:
    if a.Numa == b.Numa:
        return i32(a.Count - a.Used) < i32(b.Count - b.Used)
    return a.Numa < b.Numa


def sort_devices(devs: list) -> None:
    """sort.Sort for n <= 12 (Go 1.19 pdqsort hands such slices to insertionSort)"""
    assert len(devs) <= 12
    for i in range(1, len(devs)):
        j = i
        while j > 0 and _less(devs[j], devs[j - 1]):
            devs[j], devs[j - 1] = devs[j - 1], devs[j]
            j -= 1


def fit_in_certain_device(devices: list, request, annos: dict):
    k = SimpleNamespace(**vars(request))  # k := request (a copy)
    origin_req = k.Nums
    prevnuma = -1
    tmp_devs = []
    for i in range(len(devices) - 1, -1, -1):
        d = devices[i]
        found, numa = check_type(annos, d, k)
        if not found:
            continue
        if numa and prevnuma != d.Numa:
            k.Nums = origin_req
            prevnuma = d.Numa
            tmp_devs = []
        memreq = 0
        if d.Count <= d.Used:
            continue
        if k.Coresreq > 100:
            return False, tmp_devs
        if k.Memreq > 0:
            memreq = k.Memreq
        if k.MemPercentagereq != MEMPCT_UNSET and k.Memreq == 0:
            memreq = go_div(i32(d.Totalmem * k.MemPercentagereq), 100)
        if i32(d.Totalmem - d.Usedmem) < memreq:
            continue
        if i32(d.Totalcore - d.Usedcores) < k.Coresreq:
            continue
        if d.Totalcore == 100 and k.Coresreq == 100 and d.Used > 0:
            continue
        if d.Totalcore != 0 and d.Usedcores == d.Totalcore and k.Coresreq == 0:
            continue
        if k.Nums > 0:
            k.Nums -= 1
            tmp_devs.append(SimpleNamespace(Idx=i, UUID=d.Id, Type=k.Type, Usedmem=memreq, Usedcores=k.Coresreq))
        if k.Nums == 0:
            return True, tmp_devs
    return False, tmp_devs


def fit_in_devices(devices: list, requests: list, annos: dict):
    devs = []
    total = 0
    free = 0
    sums = 0
    for k in requests:
        sums += k.Nums
        if k.Nums > len(devices):
            return False, F32(0), devs
        sort_devices(devices)
        fit, tmp_devs = fit_in_certain_device(devices, k, annos)
        if fit:
            for val in tmp_devs:
                d = devices[val.Idx]
                total = i32(total + d.Count)
                free = i32(free + i32(d.Count - d.Used))
                d.Used = i32(d.Used + 1)
                d.Usedcores = i32(d.Usedcores + val.Usedcores)
                d.Usedmem = i32(d.Usedmem + val.Usedmem)
            devs.extend(tmp_devs)
        else:
            return False, F32(0), devs
    with np.errstate(all="ignore"):
        return True, F32(F32(total) / F32(free)) + F32(len(devices) - sums), devs


def calc_score_node(devices: list, nums: list, annos: dict):
    """the body of calcScore's loop for one node -> (in the result list?, score, devices per container)"""
    score = F32(0)
    out_devices = []
    for n in nums:
        sums = 0
        for k in n:
            sums += k.Nums
        if sums == 0:
            out_devices.append([])
            continue
        fit, nodescore, devs = fit_in_devices(devices, n, annos)
        if fit:
            out_devices.append(devs)
            with np.errstate(all="ignore"):
                score = F32(score + nodescore)
        else:
            break
    return len(out_devices) == len(nums), score, out_devices


# ---- the packed model in and out ----------------------------------------------------------------------------------------


def _node_devices(snap, row: int) -> list:
    out = []
    for s in range(int(snap.ndev[row])):
        d = DeviceUsage()
        d.Id = s  # registration slot stands for the UUID
        d.Used, d.Count, d.Usedmem, d.Totalmem, d.Usedcores, d.Totalcore, d.Numa = (int(snap.planes[f, row, s]) for f in range(7))
        d.Type = snap.type_names[int(snap.planes[7, row, s])]
        out.append(d)
    return out


def _pod_requests(batch, p: int) -> list:
    nums = []
    for c in range(int(batch.ctr_begin[p]), int(batch.ctr_begin[p + 1])):
        ctr = []
        for r in range(int(batch.req_begin[c]), int(batch.req_begin[c + 1])):
            ctr.append(SimpleNamespace(Nums=int(batch.nums[r]), Type=batch.req_type[r], Memreq=int(batch.memreq[r]),
                                       MemPercentagereq=int(batch.mempct[r]), Coresreq=int(batch.coresreq[r]), _r=r))
        nums.append(ctr)
    return nums


def score_python(snap, batch):
    """every pod against every node; the same fields as oracle.score_faithful(..., full=True)"""
    P, N = batch.n_pods, snap.n_nodes
    pb = batch.pick_begin
    best_row = np.full(P, -1, np.int32)
    score = np.zeros(P, np.float32)
    fit_count = np.zeros(P, np.int32)
    n_picks = np.zeros(P, np.int32)
    picks = np.zeros((int(pb[-1]), 3), np.int32)
    fit = np.zeros((P, N), bool)
    scores = np.zeros((P, N), np.float32)
    for p in range(P):
        nums = _pod_requests(batch, p)
        r0, r1 = batch.pod_req_range(p)
        for row in range(N):
            ok, sc, devs = calc_score_node(_node_devices(snap, row), nums, batch.annos[p])  # getNodesUsage: fresh per Filter
            if not ok:
                continue
            fit[p, row] = True
            scores[p, row] = sc
            fit_count[p] += 1
            if best_row[p] < 0 or sc >= score[p]:  # last element of the sorted list; ties: the highest row
                best_row[p], score[p] = row, sc
                n_picks[p] = int(pb[r1] - pb[r0])
                picks[int(pb[r0]):int(pb[r1])] = 0
                # one container's devices are its requests' picks in request order, Nums each
                at = int(pb[r0])
                for ctr in devs:
                    for val in ctr:
                        picks[at] = (val.UUID, val.Usedmem, val.Usedcores)
                        at += 1
    return SimpleNamespace(best_row=best_row, score=score, fit_count=fit_count, n_picks=n_picks, picks=picks, fit=fit,
                           scores=scores)
