// vgs_layout.h -- device-side data layout shared by the kernels and the C ABI.
//
// Two copies of the usage snapshot live in HBM:
//
//  raw     the 8 int32 planes exactly as the host packed them, node-major
//          (plane[f][row * dmax + slot]); source of truth, updated by deltas,
//          read by the resolve kernel.
//  walk    per node, the same devices re-ordered into the order in which the
//          reference's fitInCertainDevice visits them for the FIRST request of
//          a pod (score.go:165 sort, :92 reverse walk).  That order does not
//          depend on the pod, so it is prepared once per snapshot (K0) instead
//          of once per (pod, node).  Stored as warp tiles of 32 nodes:
//              int4 walk[tile][word][q][lane]      q = device group of 4
//          so that lane l of a warp reads node tile*32+l with 128-bit loads
//          and a warp reads 512 contiguous bytes per instruction.
#pragma once
#include <stdint.h>
#include <vector_types.h>

namespace vgs {

constexpr int kWarp = 32;

// words per device in the walk-ordered copy
enum {
    W_TOTALMEM = 0,  // Totalmem
    W_FREEMEM = 1,   // Totalmem - Usedmem      (score.go:121 left-hand side)
    W_FREECORES = 2, // Totalcore - Usedcores   (score.go:124 left-hand side)
    W_COUNT = 3,     // Count
    W_FREESLOTS = 4, // Count - Used            (score.go:47,107,170)
    W_META = 5,      // packed small fields, see M_*
    W_SCORE1 = 6,    // f32 bits of Count/(Count-Used) + (ndev-1): the node score if this
                     // device alone is picked (score.go:180 with Nums == 1)
    W_TOTALCORE = 7, // Totalcore (with W_FREECORES: Usedcores; read by the drain only)
    kWalkWords = 8
};

// W_META bits
constexpr uint32_t M_TYPE_MASK = 0x3fu;   // bits 0-5  type id
constexpr int M_NUMA_SHIFT = 6;           // bits 6-9  rank of Numa among the node's values
constexpr int M_SLOT_SHIFT = 10;          // bits 10-13 registration slot
constexpr int M_NDEV_SHIFT = 14;          // bits 14-17 len(node.Devices)
constexpr uint32_t M_NOFREE = 1u << 18;   // Count <= Used                       (score.go:107)
constexpr uint32_t M_EXCL = 1u << 19;     // Totalcore == 100 && Used > 0        (score.go:128)
constexpr uint32_t M_FULL = 1u << 20;     // Totalcore != 0 && Usedcores == Totalcore (:132)
constexpr uint32_t M_PAD = 1u << 21;      // slot >= len(node.Devices): not a device
constexpr uint32_t M_IS100 = 1u << 22;    // Totalcore == 100
constexpr uint32_t M_CORENZ = 1u << 23;   // Totalcore != 0
constexpr uint32_t M_BUSY370 = 1u << 24;  // M_T370 && Used > 0            (cambricon/device.go:97)
constexpr uint32_t M_T370 = 1u << 25;     // type flagged VGS_TYPE_IDLE_SHARE ("370" in d.Type)
static_assert(M_BUSY370 == (M_T370 >> 1), "the general kernel derives one from the other by a shift");
constexpr uint32_t M_ALWAYS = 1u << 31;   // always set (lets a request block every device)

// ordering key of a device inside one node: (numa rank, free share slots, previous position)
constexpr int kKeyPosBits = 4;
constexpr int kKeyFreeBits = 23;
constexpr int32_t kKeyFreeBias = 1 << 22;

// request as the kernels see it (built by the host from vgs_req)
struct __align__(16) DevReq {
    int32_t nums;
    int32_t memfix;     // Memreq when > 0 else 0                           (score.go:114-116)
    int32_t pct;        // MemPercentagereq, used when R_USEPCT             (score.go:117-120)
    int32_t coresreq;
    uint32_t blockmask; // W_META bits that rule a device out for this request
    uint32_t flags;     // R_*
    uint32_t mask_lo, mask_hi; // type mask
};
constexpr uint32_t R_NUMA = 1u;   // numa-bind asserted
constexpr uint32_t R_USEPCT = 2u; // memreq = Totalmem * pct / 100 per device
constexpr uint32_t R_NEWCTR = 4u; // first request of a container
constexpr uint32_t R_IDLE = 8u;   // VGS_REQ_IDLE_SHARE (K3's plain restatement; the score kernels see M_BUSY370 in blockmask)

// ordering key of the walk kernel (vgs_general.cu): [numa rank 4][free share slots + bias 21][arrival stamp 7]
constexpr int kStampBits = 7;
constexpr uint32_t kStampMask = (1u << kStampBits) - 1u;
constexpr int kFsBits = 21;
constexpr uint32_t kFsBias = 1u << 20; // |Count - Used| < VGS_MAX_SHARE = 2^20
constexpr int kNumaShift = kStampBits + kFsBits; // 28

// pod of the general (multi-request) class
struct GenPod {
    int32_t req_begin, n_reqs, orig, flags; // flags: G_*, 0 = nothing known (always correct)
};
constexpr int32_t G_UNIFORM_TYPE = 1;  // every request of the pod carries the same type mask and no R_IDLE
constexpr int32_t G_NO_USED_TRACK = 2; // no request's blockmask tests a "Used > 0" bit (M_EXCL, M_BUSY370)
constexpr int32_t G_LEGACY = 4;        // host routing only: more picks than the walk kernel's arrival stamps hold
constexpr int32_t G_SAME_FIRST = 8;    // same first request (and pod flags) as the pod before it in the class list
constexpr int32_t G_SAVE_FIRST = 16;   // the pod after it carries G_SAME_FIRST
constexpr int kDrainChunk = 64;       // pods per chunk of the batch-scale drain (= touched rows a chunk can hold)
constexpr int kDrainTop = 32;         // candidate keys kept per pod by the chunk-start pass
constexpr int kWalkMaxPicks = 100;     // sum of Nums per pod the walk kernel accepts (7-bit stamps, 12 initial)

inline int32_t gen_pod_flags(const DevReq *r, int n) {
    int32_t f = G_UNIFORM_TYPE | G_NO_USED_TRACK;
    long long picks = 0;
    for (int i = 0; i < n; i++) {
        if (r[i].mask_lo != r[0].mask_lo || r[i].mask_hi != r[0].mask_hi || (r[i].flags & R_IDLE)) f &= ~G_UNIFORM_TYPE;
        if (r[i].blockmask & (M_EXCL | M_BUSY370)) f &= ~G_NO_USED_TRACK;
        picks += r[i].nums;
    }
    if (picks > kWalkMaxPicks) f |= G_LEGACY;
    return f;
}

// Pod-sliced score kernel: per-class bit tables over the pods of one class (bit i of word
// i/32 = pod i of the class).  Row stride Wp words.  "LE" tables are cumulative: row c holds
// the pods whose threshold is among the c smallest distinct values, so a device whose
// capacity admits exactly c of the distinct values selects row c.
struct SliceTab {
    // all rows in one buffer, row r at base + r * Wp, in this order:
    //   [0, 64)            type rows: pods whose type mask has bit t
    //   [64, 64+km+1)      fixed-MB pods by Memreq (cumulative; row 0 is all zero)
    //   then kp+1 rows     percentage pods by MemPercentagereq (cumulative)
    //   then 4*(kc+1) rows all pods by Coresreq (cumulative) AND surviving the device flags:
    //                      row = cut * 4 + (excl | full << 1); excl: Coresreq == 100 pods lose
    //                      (score.go:128), full: Coresreq == 0 pods lose (:132), Coresreq > 100
    //                      pods always lose (:110)
    const uint32_t *base;
    const int32_t *memv, *pctv, *corev; // sorted distinct values
    const int32_t *orig;                // class position -> pod index of the batch
    int32_t km, kp, kc;
    int32_t n_pods, W, Wp;
    int32_t row_shift; // SliceDev rows >= VGS_MAX_TYPES are stored minus this (the score kernel's shared
                       // copy of the tables drops the type rows that cannot occur); set by the launcher
    // per-batch products of the preparation kernels (K1p), all in device memory
    uint4 *dinfo;      // [n_tiles*32*8] SliceDev, node-major, walk order
    uint8_t *pcls;     // [n_tiles*32][32] score class of each device pair (Nums == 2 class only)
    uint32_t *hset;    // [kSliceHash] open-addressing set of the distinct candidate scores
    uint32_t *gvals;   // [kSliceDict] the largest distinct scores, ascending; gvals[kSliceDict] = count
};
// what the sliced score kernel needs to know about one device, 16 bytes
struct __align__(16) SliceDev {
    uint16_t r_type, r_mem, r_pct, r_core; // table rows this device selects (r_type == r_mem: never fits)
    uint32_t a;                            // Nums == 1: score bits if picked alone; Nums == 2: Count
    uint32_t b;                            // Nums == 1: score class; Nums == 2: free slots | ndev << 24
};
constexpr int kSliceDict = 32;    // distinct candidate scores tracked exactly (the largest ones)
constexpr int kSliceHash = 4096;  // capacity of the distinct-score set (power of two)
constexpr int kSliceMaxDistinct = 1024; // distinct Memreq / Coresreq values per class
constexpr int kSliceMaxPct = 101;       // MemPercentagereq in [0, 100]
constexpr int kSliceCountPlanes = 13;   // bit-sliced fit counter: < 8192 nodes per CTA
constexpr int32_t kSliceMaxTotalmem = 21474836; // Totalmem * 100 must not wrap int32

// status word bits (device -> host)
constexpr uint32_t S_RANGE = 1u;    // snapshot value out of range
constexpr uint32_t S_TYPE = 2u;     // type id out of range
constexpr uint32_t S_NDEV = 4u;     // ndev out of range
constexpr uint32_t S_MISMATCH = 8u; // resolve kernel disagrees with score kernel
constexpr uint32_t S_DELTA = 16u;   // delta names a row/slot outside the shard
constexpr uint32_t S_BIGMEM = 32u;  // (not an error) some Totalmem outside [0, kSliceMaxTotalmem]
constexpr uint32_t S_DICTFULL = 64u; // more than kSliceHash distinct candidate scores in one class

__host__ __device__ inline uint32_t order_bits(uint32_t fbits) {
    // monotone map float bits -> unsigned, never 0 for a finite float
    return fbits ^ ((uint32_t)((int32_t)fbits >> 31) | 0x80000000u);
}
__host__ __device__ inline uint32_t unorder_bits(uint32_t ob) {
    return (ob & 0x80000000u) ? (ob ^ 0x80000000u) : ~ob;
}

struct ScoreLaunch {
    int grid, block, smem;
    int chunk_len, n_chunks, n_tile_groups, warps_per_cta;
};

} // namespace vgs
