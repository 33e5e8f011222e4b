// Shared description of the packed-weight image used by the tcgen05 kernels.
#pragma once
#include "common.cuh"
#include "tc_common.cuh"

#define TC_STAGE_BYTES 16384          // [128 n x 64 k] bf16, SW128 K-major
#define TC_SLAB_BYTES 16384           // A-operand slab [128 rows x 64 k] bf16
#define TC_VEC_BIAS 0                 // fp32 section: bias'[10][256]
#define TC_VEC_HEADW (10 * 256)       //               head weights [4][256] (density, r, g, b) with gamma folded
#define TC_VEC_HEADB (14 * 256)       //               head biases [4]
#define TC_VEC_FLOATS (14 * 256 + 4)
// Forward weights: bf16 B operand without swizzle, K-major, as a sequence of "k-cores" per (unit, n-half):
// k-core = [128 n x 8 k] = 16 n-groups x (8 rows x 16 B) = 2 KB, so any run of k-cores is one contiguous
// bulk copy and the ring stage size is a free choice (descriptor: LBO = 2 KB between k-cores, SBO = 128 B
// between n-groups).  Small stages matter: the ring's round trip (MMA hold + refill latency) bounds the
// MMA rate at bytes_in_flight / round_trip, and the hold part shrinks with the stage.
// Bias rides in the accumulator.  Every position/direction slab row carries 1.0 in columns 60..62;
// units 0, 4, 8 (which read an encoding slab anyway) hold bias' as three bf16 parts (hi + mid + lo,
// exact to ~2^-24) in those weight columns, the other units start with a two-k-core "bias block" that is
// multiplied with slab columns 48..63: k-core 0 zero, k-core 1 = (0, 0, 0, 0, hi, mid, lo, 0).
#define TC_ONES_COL 60
#define TF_KCORE_BYTES 2048
#define TF_STAGE_KCORES 8             // K = 64 per ring stage (four MMAs)
#define TF_STAGE_BYTES (TF_STAGE_KCORES * TF_KCORE_BYTES)
#define TF_BIAS_KCORES 2              // in front of the unit's first stage (same bulk copy, fifth MMA)
#define TF_SLOT_BYTES (TF_STAGE_BYTES + TF_BIAS_KCORES * TF_KCORE_BYTES)
#define TF_B_LBO 2048
#define TF_B_SBO 128
#define TC_VEC_BYTES ((TC_VEC_FLOATS * 4 + 1023) & ~1023)

// Conditional inputs (G2: states_x appended to the normalised position before the encoding, Dx = 3 + Sx):
// the states are per-RAY constants, so their share of units 0 and 4, W[:, state columns] . PE(states), is a
// per-ray bias vector [256].  The slab and the weight image keep the 60 position columns only; the state
// weights travel as an fp32 section behind the vectors ([2 units][256][20 Sx]) for the per-ray bias GEMM.
struct PackLayout {
    int Dx, Dd, EX, ED;
    int EXs;                    // encoding columns of the states (20 per state), 0 for G1
    int64_t state_w_off;        // byte offset of the fp32 state weights in the packed image
    int nk[10];                 // k-slabs (of 64) per unit
    int stage_begin[11];        // first 16 KB stage of each unit (2 stages per k-slab)
    int n_stages;
    int Kt[10];
    int64_t unit_off[10];       // forward image: byte offset of the unit's [2 n-halves][k-cores] region
    int half_kcores[10];        // k-cores per n-half: bias block (units without an encoding slab) + 8 per k-slab
    int64_t weight_bytes;
    int64_t off_W[10], off_b[10], off_gamma[10], off_beta[10];
    int64_t off_head_w[2], off_head_b[2];
};

static inline int make_pack_layout(int Dx, int Dd, PackLayout& L) {
    if (Dx < 3 || Dd < 3 || Dx > 16 || Dd > 16) return NERF_ERR_BAD_ARG;
    int64_t off[44];
    nerf_g1_param_layout(Dx, Dd, off);
    L.Dx = Dx; L.Dd = Dd; L.EX = 2 * NERF_PE_X_OCTAVES * Dx; L.ED = 2 * NERF_PE_D_OCTAVES * Dd;
    L.EXs = 2 * NERF_PE_X_OCTAVES * (Dx - 3);
    const int nkx = (2 * NERF_PE_X_OCTAVES * 3 + 63) / 64, nkd = (L.ED + 63) / 64;   // the slab holds the position columns
    int s = 0;
    for (int l = 0; l < 10; ++l) {
        L.nk[l] = (l == 0) ? nkx : (l == 4 ? 4 + nkx : (l == 8 ? 4 + nkd : 4));
        L.Kt[l] = (l == 0) ? L.EX : (l == 4 ? 256 + L.EX : (l == 8 ? 256 + L.ED : 256));
        L.stage_begin[l] = s; s += 2 * L.nk[l];
        L.off_W[l] = off[4 + 4 * l]; L.off_b[l] = off[5 + 4 * l];
        L.off_gamma[l] = off[6 + 4 * l]; L.off_beta[l] = off[7 + 4 * l];
    }
    L.stage_begin[10] = s; L.n_stages = s;
    int64_t wo = 0;
    for (int l = 0; l < 10; ++l) {
        const bool blk = (l != 0 && l != 4 && l != 8);
        L.half_kcores[l] = (blk ? TF_BIAS_KCORES : 0) + 8 * L.nk[l];
        L.unit_off[l] = wo; wo += 2ll * L.half_kcores[l] * TF_KCORE_BYTES;
    }
    L.weight_bytes = wo;
    L.off_head_w[0] = off[0]; L.off_head_b[0] = off[1]; L.off_head_w[1] = off[2]; L.off_head_b[1] = off[3];
    return NERF_OK;
}

__host__ __device__ static inline bool tc_unit_has_bias_block(int l) { return l != 0 && l != 4 && l != 8; }
static inline int64_t tc_packed_vec_offset(const PackLayout& L) { return L.weight_bytes; }
static inline int64_t tc_packed_state_w_offset(const PackLayout& L) { return L.weight_bytes + TC_VEC_BYTES; }
static inline int64_t tc_packed_total(const PackLayout& L) { return L.weight_bytes + TC_VEC_BYTES + ((2ll * 256 * L.EXs * 4 + 1023) & ~1023ll); }

// ---- fp32-grade forward ("bf16x3", fused_tc_x3.cu): every operand is split a = a_hi + a_lo (two bf16), and
// a w ~ a_hi w_hi + a_lo w_hi + a_hi w_lo (fp32 accumulate; the dropped lo*lo term is ~2^-18 relative).  Weight image:
// per (unit, n-half) [bias block, 4 KB, units without an encoding slab] + 2 nk groups of K = 32:
// [4 k-cores hi | 4 k-cores lo] = 16 KB, one contiguous bulk copy per ring stage.  Dx = Dd = 3 only.
#define X3_GROUP_BYTES 16384
struct X3Layout { int nk[10]; int64_t unit_off[10]; int64_t half_bytes[10]; int64_t weight_bytes; };
static inline void make_x3_layout(X3Layout& X) {
    PackLayout L; make_pack_layout(3, 3, L);
    int64_t o = 0;
    for (int l = 0; l < 10; ++l) {
        X.nk[l] = L.nk[l];
        X.half_bytes[l] = (tc_unit_has_bias_block(l) ? TF_BIAS_KCORES * TF_KCORE_BYTES : 0) + 2ll * L.nk[l] * X3_GROUP_BYTES;
        X.unit_off[l] = o; o += 2 * X.half_bytes[l];
    }
    X.weight_bytes = o;
}
static inline int64_t tc_packed_x3_total(const X3Layout& X) { return X.weight_bytes + TC_VEC_BYTES; }

// ---- workspace of the bf16 training step (tc_train.cu): byte offsets, all 1024-aligned ----
#define TC_GW_LD 336                   // dW' row: 256 hidden + 64 encoding + 16 x (column sum of dz = db')
struct TrainLayout {
    int64_t P, P_pad; int n_tiles, tr_tiles;
    int64_t off_pex, off_ped, off_act, off_dz, off_stats, off_dh, off_gw, off_gh, off_hsum, total;
    int64_t zero_begin, zero_bytes;    // [gw | gh | hsum | gs]: accumulated with atomics, cleared every step
    // conditional inputs (EXs > 0): PE(states) [B][EXs], per-ray bias [2][B][256], per-ray sums of dz [2][B][256],
    // state-weight gradients [2][256][EXs] (all fp32)
    int EXs; int64_t off_pes, off_rb, off_rsum, off_gs;
};
static inline void make_train_layout(int64_t B, int S, TrainLayout& T, int Dx = 3) {
    auto up = [](int64_t x) { return (x + 1023) & ~(int64_t)1023; };
    T.P = B * S; T.n_tiles = (int)((T.P + 127) / 128); T.tr_tiles = (T.n_tiles + 1) & ~1; T.P_pad = (int64_t)T.tr_tiles * 128;
    int64_t o = 0;
    T.off_pex = o; o += (int64_t)T.tr_tiles * TC_SLAB_BYTES;
    T.off_ped = o; o += (int64_t)T.tr_tiles * TC_SLAB_BYTES;
    T.off_act = o; o += 10ll * T.tr_tiles * 4 * TC_SLAB_BYTES;
    T.off_dz = o; o += 10ll * T.tr_tiles * 4 * TC_SLAB_BYTES;
    T.off_stats = o; o += up(10ll * T.P_pad * 8);
    T.off_dh = o; o += (int64_t)T.tr_tiles * 4096;
    T.off_gw = o; T.zero_begin = o; o += up(10ll * 256 * TC_GW_LD * 4);
    T.off_gh = o; o += up(2ll * 256 * 16 * 4);
    T.off_hsum = o; o += 1024;
    T.EXs = 2 * NERF_PE_X_OCTAVES * (Dx - 3);
    T.off_gs = o; o += up(2ll * 256 * T.EXs * 4);
    T.zero_bytes = o - T.zero_begin;
    T.off_pes = o; o += up(B * (int64_t)T.EXs * 4);
    T.off_rb = o; o += T.EXs ? up(2ll * B * 256 * 4) : 0;
    T.off_rsum = o; o += T.EXs ? up(2ll * B * 256 * 4) : 0;
    T.total = o;
}

// ---- workspace of the fp32-grade ("bf16x3") training step (tc_train_x3.cu): the bf16 step's layout holds the hi parts,
// a second set of images behind it the lo parts (activations, dz, the two encodings, the (d_sigma, d_rgb) image)
// The (hi, lo) slab images of the bf16x3 step in HBM.  In shared memory a tile's image is 4 slabs (hidden) or 1 slab (encoding)
// of [128 points x 128 B], hi and lo in separate buffers; in HBM the two parts are interleaved in 4 KB pieces of 32 points:
//     hidden    [unit][tile][quarter q = point / 32][slab s][part hi|lo][4096 B]        128 KB per (unit, tile)
//     encoding        [tile][quarter q][part][4096 B]                                    32 KB per tile
//     small operand   [tile][quarter q][part][1024 B]    ((d_sigma, d_rgb) rows: 8 points x 2 column halves x 16 B)
// so that the weight-gradient GEMM, whose contraction runs over the points, fetches a 32-point stage of one CTA (2 slabs x
// 2 parts) as ONE 16 KB bulk copy.  The forward and data-gradient kernels move the images as 16-byte thread chunks and apply
// the permutation for free.
#define X3_IMG_BYTES (8 * TC_SLAB_BYTES)      // hidden image of one (unit, tile), both parts
#define X3_ENC_IMG_BYTES (2 * TC_SLAB_BYTES)
#define X3_DH_IMG_BYTES 8192
// byte offset inside a hidden image of the 16-byte chunk at offset o (o < 65536) of the shared-memory part `part`
__host__ __device__ static inline uint32_t x3_img_offset(uint32_t o, uint32_t part) {
    return ((((o >> 12) & 3u) * 4u + (o >> 14)) * 2u + part) * 4096u + (o & 4095u);
}
// ... inside an encoding image (o < 16384)
__host__ __device__ static inline uint32_t x3_enc_img_offset(uint32_t o, uint32_t part) { return ((o >> 12) * 2u + part) * 4096u + (o & 4095u); }
// ... of point row `row`'s 16-byte entry (column half `ch`) inside a small-operand image
__host__ __device__ static inline uint32_t x3_dh_img_offset(uint32_t row, uint32_t ch, uint32_t part) {
    return ((row >> 5) * 2u + part) * 1024u + ((row >> 3) & 3u) * 256u + ch * 128u + (row & 7u) * 16u;
}

struct X3TrainLayout {
    TrainLayout T;       // the fp32 outputs, statistics and flags of the bf16 step; its slab-image regions are re-cut below
    int64_t off_act3, off_dz3;     // [10][tr_tiles] hidden images
    int64_t off_pex3, off_ped3;    // [tr_tiles] encoding images
    int64_t off_dh3;               // [tr_tiles] small-operand images
    int64_t off_mask;    // [10][P_pad][8] uint32: the ReLU mask of every unit, one bit per column (bit c & 31 of word c >> 5)
    int64_t total;
};
static inline void make_x3_train_layout(int64_t B, int S, X3TrainLayout& X) {
    make_train_layout(B, S, X.T, 3);
    int64_t o = X.T.total;
    const int64_t unit_images = 10ll * X.T.tr_tiles * X3_IMG_BYTES;
    X.off_act3 = X.T.off_act;                       // T's act and dz regions are adjacent: together one x3 region
    X.off_dz3 = o; o += unit_images;
    X.off_pex3 = X.T.off_pex;                       // likewise T's pex and ped
    X.off_ped3 = o; o += (int64_t)X.T.tr_tiles * X3_ENC_IMG_BYTES;
    X.off_dh3 = o; o += (int64_t)X.T.tr_tiles * X3_DH_IMG_BYTES;
    X.off_mask = o; o += 10ll * X.T.P_pad * 32;
    X.total = o;
}

