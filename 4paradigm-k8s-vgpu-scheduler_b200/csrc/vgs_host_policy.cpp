// vgs_host_policy.cpp -- host-side string policy and wire format.
//
// In the Go extender these stay in Go: the shim calls the reference's own checkType once
// per distinct device type string and packs the answers into vgs_req.typemask.  This C++
// mirror exists so that hosts without Go (this repo's C++ Filter mirror, the tests, the
// bench) build exactly the same masks.
//
// Mirrors: pkg/scheduler/score.go:71-84 (checkType), pkg/device/nvidia/device.go:62-94
// (checkGPUtype), :96-105 (assertNuma), :107-112 (CheckType), pkg/util/util.go:110-126
// (EncodeContainerDevices / EncodePodDevices).
#include <algorithm>
#include <cstdint>
#include <string>
#include <vector>

#include "../../include/vgpuscore.h"

namespace {

// Go strings.ToUpper on ASCII; bytes >= 0x80 pass through unchanged (Go maps them by
// Unicode case tables; device model strings are ASCII).
std::string upper(const std::string &s) {
    std::string r(s);
    for (char &ch : r)
        if (ch >= 'a' && ch <= 'z') ch = (char)(ch - 'a' + 'A');
    return r;
}

bool contains(const std::string &hay, const std::string &needle) {
    return hay.find(needle) != std::string::npos; // true for the empty needle, like Go
}

// strings.Split(list, ","): every token, empty ones included
std::vector<std::string> split_commas(const std::string &list) {
    std::vector<std::string> out;
    size_t start = 0;
    for (;;) {
        size_t comma = list.find(',', start);
        if (comma == std::string::npos) {
            out.push_back(list.substr(start));
            return out;
        }
        out.push_back(list.substr(start, comma - start));
        start = comma + 1;
    }
}

// is any token of the annotation value a case-insensitive substring of the card type?
bool names_card(const std::string &anno, const std::string &card_upper) {
    for (const std::string &tok : split_commas(anno)) // one token when there is no comma
        if (contains(card_upper, upper(tok))) return true;
    return false;
}

// checkGPUtype: use-gputype wins outright; nouse-gputype only when use is absent
bool gpu_type_allowed(const char *use, const char *nouse, const std::string &card) {
    const std::string cu = upper(card);
    if (use) return names_card(use, cu);
    if (nouse) return !names_card(nouse, cu);
    return true;
}

} // namespace

extern "C" {

uint64_t vgs_host_typemask(const char *req_type, const char *use_gputype, const char *nouse_gputype,
                           const char *const *types, int32_t n_types) {
    uint64_t mask = 0;
    if (!req_type || !types) return 0;
    const std::string rt(req_type);
    if (n_types > VGS_MAX_TYPES) n_types = VGS_MAX_TYPES;
    for (int32_t t = 0; t < n_types; t++) {
        if (!types[t]) continue;
        const std::string card(types[t]);
        if (!contains(card, rt)) continue; // score.go:73  d.Type must contain the request type
        // vendor lookup (score.go:76-81): only the NVIDIA policy is implemented; any other
        // request type is "Unrecognized device" -> no device passes.
        if (rt != "NVIDIA") continue;
        if (gpu_type_allowed(use_gputype, nouse_gputype, card)) mask |= 1ull << t;
    }
    return mask;
}

int32_t vgs_host_numa_bind(const char *v) {
    if (!v) return 0;
    // strconv.ParseBool's accepted spellings of true
    static const char *const yes[] = {"1", "t", "T", "TRUE", "true", "True"};
    const std::string s(v);
    for (const char *y : yes)
        if (s == y) return 1;
    return 0;
}

int32_t vgs_host_encode_pod_devices(int32_t n_ctrs, const int32_t *picks_per_ctr, const char *const *uuids,
                                    const char *const *types, const vgs_pick *picks, char *out, int32_t cap) {
    if (n_ctrs < 0 || (n_ctrs > 0 && !picks_per_ctr) || !out || cap < 1) return VGS_EINVAL;
    std::string s;
    int32_t t = 0;
    for (int32_t c = 0; c < n_ctrs; c++) {
        if (c) s += ';';
        for (int32_t j = 0; j < picks_per_ctr[c]; j++, t++) {
            s += uuids[t];
            s += ',';
            s += types[t];
            s += ',';
            s += std::to_string(picks[t].usedmem);
            s += ',';
            s += std::to_string(picks[t].usedcores);
            s += ':';
        }
    }
    if ((int64_t)s.size() + 1 > cap) return VGS_ETOOBIG;
    std::copy(s.begin(), s.end(), out);
    out[s.size()] = 0;
    return (int32_t)s.size();
}

} // extern "C"
