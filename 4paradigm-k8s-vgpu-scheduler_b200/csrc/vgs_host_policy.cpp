// vgs_host_policy.cpp -- host-side string policy and wire format.
//
// In the Go extender these stay in Go: the shim calls the reference's own checkType once
// per distinct device type string and packs the answers into vgs_req.typemask.  This C++
// mirror exists so that hosts without Go (this repo's C++ Filter mirror, the tests, the
// bench) build exactly the same masks.
//
// Mirrors: pkg/scheduler/score.go:71-84 (checkType), pkg/device/nvidia/device.go:62-94
// (checkGPUtype), :96-105 (assertNuma), :107-112 (CheckType), pkg/util/util.go:110-126
// (EncodeContainerDevices / EncodePodDevices); strings.ToUpper / strings.Map / unicode.ToUpper of Go 1.19.
#include <algorithm>
#include <cstdint>
#include <string>
#include <vector>

#include "../../include/vgpuscore.h"

namespace {

// ---- Go 1.19 strings.ToUpper ---------------------------------------------------------------------------------
// ASCII strings are folded byte by byte; anything else goes rune by rune through unicode.ToUpper = the SIMPLE
// uppercase mapping of Unicode 13.0.0 (one rune to one rune: U+00DF stays, U+0131 and U+017F become ASCII I and S).
// strings.Map decodes like a range loop, so every byte that is not part of a valid encoding becomes U+FFFD in the
// result.  The table is generated (tools/make_upper_table.py).
struct UpperRun {
    uint32_t lo, hi;
    int32_t stride, delta;
};
const UpperRun kUpperRuns[] = {
#include "vgs_upper_table.inc"
};
constexpr uint32_t kRuneError = 0xFFFD;

uint32_t rune_upper(uint32_t r) {
    if (r < 0x80) return (r >= 'a' && r <= 'z') ? r - 32 : r;
    const UpperRun *b = kUpperRuns, *e = kUpperRuns + sizeof kUpperRuns / sizeof kUpperRuns[0];
    const UpperRun *it = std::upper_bound(b, e, r, [](uint32_t x, const UpperRun &u) { return x < u.lo; });
    if (it == b) return r;
    --it; // the last run that starts at or below r
    if (r <= it->hi && (r - it->lo) % (uint32_t)it->stride == 0) return (uint32_t)((int32_t)r + it->delta);
    return r;
}

// utf8.DecodeRuneInString: (rune, width); (U+FFFD, 1) for a byte that does not start a valid encoding (overlong forms,
// surrogates, values above U+10FFFF and truncated sequences included)
uint32_t decode_rune(const std::string &s, size_t i, size_t *width) {
    const size_t n = s.size() - i;
    auto at = [&](size_t k) { return (uint32_t)(unsigned char)s[i + k]; };
    auto cont = [&](size_t k, uint32_t lo, uint32_t hi) { return k < n && at(k) >= lo && at(k) <= hi; };
    const uint32_t b0 = at(0);
    *width = 1;
    if (b0 < 0x80) return b0;
    if (b0 >= 0xC2 && b0 <= 0xDF) {
        if (!cont(1, 0x80, 0xBF)) return kRuneError;
        *width = 2;
        return ((b0 & 0x1F) << 6) | (at(1) & 0x3F);
    }
    if (b0 >= 0xE0 && b0 <= 0xEF) {
        const uint32_t lo = b0 == 0xE0 ? 0xA0 : 0x80, hi = b0 == 0xED ? 0x9F : 0xBF;
        if (!cont(1, lo, hi) || !cont(2, 0x80, 0xBF)) return kRuneError;
        *width = 3;
        return ((b0 & 0x0F) << 12) | ((at(1) & 0x3F) << 6) | (at(2) & 0x3F);
    }
    if (b0 >= 0xF0 && b0 <= 0xF4) {
        const uint32_t lo = b0 == 0xF0 ? 0x90 : 0x80, hi = b0 == 0xF4 ? 0x8F : 0xBF;
        if (!cont(1, lo, hi) || !cont(2, 0x80, 0xBF) || !cont(3, 0x80, 0xBF)) return kRuneError;
        *width = 4;
        return ((b0 & 0x07) << 18) | ((at(1) & 0x3F) << 12) | ((at(2) & 0x3F) << 6) | (at(3) & 0x3F);
    }
    return kRuneError;
}

void append_rune(std::string &out, uint32_t r) {
    if (r < 0x80) {
        out.push_back((char)r);
    } else if (r < 0x800) {
        out.push_back((char)(0xC0 | (r >> 6)));
        out.push_back((char)(0x80 | (r & 0x3F)));
    } else if (r < 0x10000) {
        out.push_back((char)(0xE0 | (r >> 12)));
        out.push_back((char)(0x80 | ((r >> 6) & 0x3F)));
        out.push_back((char)(0x80 | (r & 0x3F)));
    } else {
        out.push_back((char)(0xF0 | (r >> 18)));
        out.push_back((char)(0x80 | ((r >> 12) & 0x3F)));
        out.push_back((char)(0x80 | ((r >> 6) & 0x3F)));
        out.push_back((char)(0x80 | (r & 0x3F)));
    }
}

std::string upper(const std::string &s) {
    bool ascii = true;
    for (char ch : s) ascii = ascii && (unsigned char)ch < 0x80;
    std::string r;
    if (ascii) {
        r = s;
        for (char &ch : r)
            if (ch >= 'a' && ch <= 'z') ch = (char)(ch - 'a' + 'A');
        return r;
    }
    r.reserve(s.size() + 4);
    for (size_t i = 0; i < s.size();) {
        size_t w;
        const uint32_t c = decode_rune(s, i, &w);
        append_rune(r, rune_upper(c));
        i += w;
    }
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

// strconv.Atoi with the error ignored, as util.DecodeNodeDevices does (`count, _ := ...`):
// a syntax error yields 0, a range error the clamped int64; then Go converts with int32(x).
int32_t atoi_go_i32(const std::string &t) {
    size_t i = 0;
    bool neg = false;
    if (i < t.size() && (t[i] == '+' || t[i] == '-')) neg = t[i++] == '-';
    if (i >= t.size()) return 0;
    unsigned long long acc = 0;
    bool over = false;
    for (; i < t.size() && !over; i++) { // ParseUint stops at the first overflow, before any later bad byte
        if (t[i] < '0' || t[i] > '9') return 0;
        const unsigned long long d = (unsigned long long)(t[i] - '0');
        if (acc > (0xffffffffffffffffull - d) / 10) over = true;
        else acc = acc * 10 + d;
    }
    long long v;
    if (over || acc > (neg ? 0x8000000000000000ull : 0x7fffffffffffffffull))
        v = neg ? (long long)0x8000000000000000ull : 0x7fffffffffffffffll;
    else
        v = neg ? -(long long)acc : (long long)acc;
    return (int32_t)(uint32_t)(unsigned long long)v;
}

// strconv.ParseInt(t, 10, 32) with the error ignored: 0 on a syntax error, the nearest int32
// limit on a range error (util.go:150-153)
int32_t parse_int32_go(const std::string &t) {
    size_t i = 0;
    bool neg = false;
    if (i < t.size() && (t[i] == '+' || t[i] == '-')) neg = t[i++] == '-';
    if (i >= t.size()) return 0;
    unsigned long long acc = 0;
    for (; i < t.size(); i++) {
        if (t[i] < '0' || t[i] > '9') return 0;
        acc = acc * 10 + (unsigned long long)(t[i] - '0');
        if (acc > 0xffffffffull) return neg ? INT32_MIN : INT32_MAX; // ParseUint(.., 32) stops here with ErrRange
    }
    if (neg) return acc > 0x80000000ull ? INT32_MIN : (int32_t)(-(long long)acc);
    return acc > 0x7fffffffull ? INT32_MAX : (int32_t)acc;
}

} // namespace

extern "C" {

int32_t vgs_host_to_upper(const char *s, char *out, int32_t cap) {
    if (!s) return -1;
    const std::string u = upper(std::string(s));
    if (out && cap > 0) {
        const size_t n = std::min(u.size(), (size_t)cap - 1);
        std::copy(u.begin(), u.begin() + n, out);
        out[n] = 0;
    }
    return (int32_t)u.size();
}

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
        // vendor lookup (score.go:76-81).  NVIDIA (nvidia/device.go:107-112) and Hygon DCU
        // (hygon/device.go:79-84) share the same use/nouse list logic -- the caller passes the
        // vendor's own annotation values (nvidia.com/{use,nouse}-gputype or
        // hygon.com/{use,nouse}-dcutype).  Cambricon's CheckType depends on d.Used and
        // n.Memreq: vgs_host_typemask_mlu below.  Any other type is "Unrecognized device".
        if (rt != "NVIDIA" && rt != "DCU") continue;
        if (gpu_type_allowed(use_gputype, nouse_gputype, card)) mask |= 1ull << t;
    }
    return mask;
}

uint64_t vgs_host_typemask_mlu(int32_t memreq, const char *use_mlutype, const char *nouse_mlutype,
                               const char *const *types, int32_t n_types, uint32_t *req_flags) {
    if (req_flags) *req_flags = memreq == 0 ? VGS_REQ_IDLE_SHARE : 0u; // cambricon/device.go:97-99
    uint64_t mask = 0;
    if (!types) return 0;
    if (n_types > VGS_MAX_TYPES) n_types = VGS_MAX_TYPES;
    for (int32_t t = 0; t < n_types; t++) {
        if (!types[t]) continue;
        const std::string card(types[t]);
        if (!contains(card, "MLU")) continue;                     // score.go:73 with n.Type == "MLU"
        if (!contains(card, "370") && memreq != 0) continue;      // cambricon/device.go:94-96
        if (gpu_type_allowed(use_mlutype, nouse_mlutype, card)) mask |= 1ull << t; // checkMLUtype :60-91
    }
    return mask;
}

uint32_t vgs_host_type_flags(const char *type_name) {
    if (!type_name) return 0;
    return contains(std::string(type_name), "370") ? VGS_TYPE_IDLE_SHARE : 0u;
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

int32_t vgs_host_decode_node_devices(const char *str, int32_t max_dev, int32_t *count, int32_t *devmem,
                                     int32_t *devcore, int32_t *numa, int32_t *health, int32_t *id_off,
                                     int32_t *id_len, int32_t *type_off, int32_t *type_len) {
    if (!str || max_dev < 0) return VGS_EINVAL;
    const std::string s(str);
    if (s.find(':') == std::string::npos) return VGS_EINVAL; // util.go:69-71
    int32_t n = 0;
    size_t start = 0;
    for (;;) { // strings.Split(str, ":")
        size_t colon = s.find(':', start);
        const size_t end = colon == std::string::npos ? s.size() : colon;
        const std::string val = s.substr(start, end - start);
        if (val.find(',') != std::string::npos) { // :75
            std::vector<size_t> cut; // item boundaries inside val
            cut.push_back(0);
            for (size_t i = 0; i < val.size(); i++)
                if (val[i] == ',') cut.push_back(i + 1);
            if (cut.size() != 7) return VGS_EINVAL; // :77, :93-95
            auto item = [&](int k) {
                const size_t b = cut[k], e = k + 1 < 7 ? cut[k + 1] - 1 : val.size();
                return val.substr(b, e - b);
            };
            if (n >= max_dev) return VGS_ETOOBIG;
            count[n] = atoi_go_i32(item(1));
            devmem[n] = atoi_go_i32(item(2));
            devcore[n] = atoi_go_i32(item(3));
            numa[n] = atoi_go_i32(item(5));
            health[n] = vgs_host_numa_bind(item(6).c_str()); // strconv.ParseBool, error -> false
            id_off[n] = (int32_t)(start + cut[0]), id_len[n] = (int32_t)item(0).size();
            type_off[n] = (int32_t)(start + cut[4]), type_len[n] = (int32_t)item(4).size();
            n++;
        }
        if (colon == std::string::npos) break;
        start = colon + 1;
    }
    return n;
}

int32_t vgs_host_decode_pod_devices(const char *str, int32_t max_ctrs, int32_t max_dev, int32_t *ctr_n,
                                    int32_t *uuid_off, int32_t *uuid_len, int32_t *type_off, int32_t *type_len,
                                    int32_t *usedmem, int32_t *usedcores) {
    if (!str || max_ctrs < 0 || max_dev < 0) return VGS_EINVAL;
    const std::string s(str);
    if (s.empty()) return 0; // util.go:159-161
    int32_t nc = 0, nd = 0;
    size_t cstart = 0;
    for (;;) { // strings.Split(str, ";")
        const size_t semi = s.find(';', cstart);
        const size_t cend = semi == std::string::npos ? s.size() : semi;
        if (nc >= max_ctrs) return VGS_ETOOBIG;
        ctr_n[nc] = 0;
        size_t start = cstart; // DecodeContainerDevices (:128-156) on s[cstart, cend); empty -> no devices
        while (cend > cstart) {  // strings.Split(piece, ":")
            size_t colon = s.find(':', start);
            if (colon == std::string::npos || colon > cend) colon = cend;
            const std::string val = s.substr(start, colon - start);
            if (val.find(',') != std::string::npos) { // :141
                std::vector<size_t> cut{0};
                for (size_t i = 0; i < val.size(); i++)
                    if (val[i] == ',') cut.push_back(i + 1);
                if (cut.size() < 4) return 0; // :144-146 error, swallowed by DecodePodDevices (:167-169)
                auto len = [&](int k) { return (k + 1 < (int)cut.size() ? cut[k + 1] - 1 : val.size()) - cut[k]; };
                if (nd >= max_dev) return VGS_ETOOBIG;
                uuid_off[nd] = (int32_t)(start + cut[0]), uuid_len[nd] = (int32_t)len(0);
                type_off[nd] = (int32_t)(start + cut[1]), type_len[nd] = (int32_t)len(1);
                usedmem[nd] = parse_int32_go(val.substr(cut[2], len(2)));
                usedcores[nd] = parse_int32_go(val.substr(cut[3], len(3)));
                nd++;
                ctr_n[nc]++;
            }
            if (colon >= cend) break;
            start = colon + 1;
        }
        nc++;
        if (semi == std::string::npos) break;
        cstart = semi + 1;
    }
    return nc;
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
