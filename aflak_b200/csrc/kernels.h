// kernels.h -- internal C++ interface between the host layer (host.cpp) and the
// sm_100a kernels (kernels.cu).  Device pointers + a stream in, no ownership.
// Not part of the public ABI (that is include/aflak_b200.h).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace afk {

// ---- axis-0 windowed reduction (K1/K2 of DESIGN.md) -------------------------
constexpr int kMaxWindows = 3;
constexpr int kReduceKGroups = 8;   // k-groups (one warp each) per CTA in the small-image variant of the reduce kernel
constexpr int kReduceKgCtasPerSm = 2; // resident CTAs per SM of that variant (<= 128 registers: no spills)
constexpr int kMaxPeers = 8;        // GPUs of one box (NVSwitch domain) a result band can be written to
constexpr int kReduceBlock = 256;   // threads per CTA
constexpr int kReduceCtasPerSm = 3; // resident CTAs per SM the launch is sized for (<= 85 registers per thread)

enum ReduceEpilogue : int {
    EPI_SUM = 0,        // out = S0                                   (integral, lib.rs:1324)
    EPI_MEAN = 1,       // out = S0 / n0                              (average,  lib.rs:1328)
    EPI_MEAN_RATIO = 2, // out = (S0/n0) / (S1/n1)                    (average, average, ratio: BASELINE C3)
    EPI_EW = 3,         // on=S0/n0 off1=S1/n1 off2=S2/n2; off=a*off1+b*off2; EW(off,on,fl,max,sign)  (templates.rs:143-463)
};

enum SumMode : int {
    SUM_INHERIT = -1,    // host layer only: no per-thread override, use the process default
    SUM_COMPENSATED = 0, // Kahan per thread, f64 combine of split-K partials (<= 1e-5 of the f64 oracle)
    SUM_SEQUENTIAL = 1,  // plain left-to-right f32 in k order, never split: bit-exact with ndarray sum_axis
};

struct ReduceArgs {
    const float *cube; // [nz][n2] frames, n2 contiguous
    int64_t n2;        // pixels per frame == frame stride in elements
    int nwin;
    int64_t start[kMaxWindows];
    int64_t end[kMaxWindows]; // exclusive
    int epilogue;
    int mode;
    float p[4];        // EPI_EW: a, b, fl, max ; sign in ew_sign
    float ew_sign;     // -1 emission, +1 absorption
    float *out;        // [n2] (may be a peer-mapped pointer)
    unsigned *minmax;  // optional: {~key(min), key(max), any-non-NaN} accumulated with atomicMax, or nullptr
    unsigned *range_out; // optional: the same three words (+ a zero) for the result image, written once by the last CTA -- needs
                         // no zeroing beforehand (the range memo kept beside an image's pixels)
    // fused gather: the same result is also stored to n_extra_out more buffers (peer-mapped over NVLink),
    // and when the whole grid has stored, `epoch` is written to each signal word (release, system scope).
    int n_extra_out;
    float *extra_out[kMaxPeers - 1];
    int n_signal;
    unsigned *signal[kMaxPeers];
    unsigned *done;    // device counter, zero-initialised, self-resetting; required when n_signal > 0
    unsigned epoch;
    // sharded min/max exchange: with range_out set, the last CTA also copies the band's four range words to
    // n_range_peer more places (this rank's entry of every peer's range table) before it signals
    int n_range_peer;
    unsigned *range_peer[kMaxPeers - 1];
    // NVLink-SHARP multicast: `out` is a multicast mapping (one multimem.st reaches every GPU's copy, this one included);
    // n_extra_out must be 0 then.  Flags and range entries stay unicast.
    int multicast_out;
};

struct ReduceWorkspace {
    void *partials;       // device scratch, >= reduce_workspace_bytes()
    size_t partial_bytes;
    unsigned *counters;   // device, zero-initialised, >= reduce_counter_count() entries, self-resetting
    size_t n_counters;
};

struct ReducePlan {
    int vec;             // 4 (float4 lanes) or 1
    int kg;              // k-groups per CTA: 1, or kReduceKGroups for small images (tile = 256/kg lanes)
    int64_t ntiles;
    int64_t ktot;
    int64_t units;       // ntiles * ktot
    int64_t q;           // units per CTA
    int grid;
    size_t partial_bytes;
    size_t n_counters;
    size_t range_slot_offset;   // byte offset of the per-CTA range slots inside the partials scratch
    size_t range_counter_index; // index of the range memo's arrival counter
    // small images in compensated mode: reduce_cluster_kernel (thread-block clusters, merge through distributed shared
    // memory).  cluster == 0: the stream-K reduce_kernel above.
    int cluster;         // CTAs per cluster (k split), 0 = not used
    int tile4;           // float4 lanes per pixel tile
    int jsub;            // interleaved frame sub-groups per CTA
    int cl_unroll;       // independent 16-byte loads per thread
    size_t cl_smem;      // dynamic shared memory per CTA
};

ReducePlan reduce_plan(const ReduceArgs &a, int sm_count);
struct ReduceVariant { int vec, comp, nw, unroll, min_blocks, kg; }; // template arguments reduce_launch instantiates
ReduceVariant reduce_variant(const ReduceArgs &a, const ReducePlan &plan);
// the kernel instantiation and launch shape of this plan as text (reports name the kernel that ran)
void reduce_describe(const ReduceArgs &a, const ReducePlan &plan, char *buf, size_t cap);
cudaError_t reduce_launch(const ReduceArgs &a, const ReducePlan &plan, const ReduceWorkspace &ws, cudaStream_t s);

// stream-ordered wait until flags[i] >= epoch for i < n (n <= kMaxPeers).  A flag that times out (bound derived from the
// device's clock rate) sets *err to 1+i and band i of `image` is overwritten with NaN.
struct GatherWaitArgs {
    const unsigned *flags;
    int n;
    unsigned epoch;
    unsigned *err;
    float *image;               // this rank's copy of the full image of that epoch
    int64_t band_off[kMaxPeers]; // element offset / length of rank i's band in it
    int64_t band_len[kMaxPeers];
    // min/max exchange: ranges[i*4 ..] = {~key(min), key(max), any, 0} of band i (arrived with its flag); their combination
    // is written to combined[0..3] -- the range memo of the gathered image.  Both nullptr when not used.
    const unsigned *ranges;
    unsigned *combined;
};
cudaError_t gather_wait_launch(const GatherWaitArgs &a, double timeout_s, int device, cudaStream_t s);

// Band broadcast (RGBA bands of the sharded display chain): copy n16 16-byte words of `src` to ndst peer-mapped
// destinations, then publish `epoch` to nsig flag words (release, system scope) once the whole grid has stored.
struct BandBroadcastArgs {
    const uint4 *src;
    int64_t n16;
    int ndst;
    uint4 *dst[kMaxPeers - 1];
    unsigned *done; // zero-initialised, self-resetting arrival counter
    int nsig;
    unsigned *sig[kMaxPeers];
    unsigned epoch;
    uint4 *mc_dst;  // NVLink-SHARP multicast mapping of the destination (then ndst == 0 and src is the local copy), or nullptr
};
cudaError_t band_broadcast_launch(const BandBroadcastArgs &a, int sm_count, cudaStream_t s);

// ---- other axis-0 reductions (axis0.cu) --------------------------------------------------------
enum Axis0Op : int {
    A0_VARIANCE = 0,      // var_axis(Axis(0), 1.0)                         lib.rs:1332-1334
    A0_STDDEV = 1,        // std_axis(Axis(0), 1.0)                         lib.rs:1336-1338
    A0_ARGFOLD = 2,       // run_create_argmap + run_minmax in one pass     lib.rs:1374-1464
    A0_CENTROID = 3,      // run_centroid                                   lib.rs:1466-1527
    A0_CENTROID_MASK = 4, // run_centroid_with_mask                         lib.rs:1529-1582
    A0_MEDIAN_OP = 5,     // run_median (median_launch)                     lib.rs:1340-1372
};

struct Axis0Args {
    int op;
    int mode;          // SumMode: sequential = the reference's own f32 order (bit-exact), default = f64 accumulators
    const float *cube; // [nz][n2]
    int64_t n2, nz;
    int64_t start, end;
    int is_min;
    int64_t r1, r2;    // A0_ARGFOLD: offsets added to the winning index for out / out2
    int out2_kind;     // A0_ARGFOLD: 0 none, 1 max/min fold (run_minmax), 2 second arg map
    int use_wcs;       // world(p) = (f32)(crval + cdelt * (p + 1 - crpix)) in f64; otherwise world(p) = p
    double crval, crpix, cdelt;
    const float *mask_lo, *mask_hi; // A0_CENTROID_MASK: per-pixel [start, end)
    float *out, *out2;
};

struct Axis0Plan {
    int vec;
    int block; // threads per CTA
    int64_t ntiles;
    int ksplit;
    int64_t chunk;
    size_t partial_bytes;
};
Axis0Plan axis0_plan(const Axis0Args &a, int sm_count);
// partials: >= plan.partial_bytes of device scratch (may be nullptr when 0); flag: one zero-initialised word that is
// set where the reference would panic (A0_CENTROID_MASK ranges outside the cube)
cudaError_t axis0_launch(const Axis0Args &a, const Axis0Plan &plan, void *partials, unsigned *flag, cudaStream_t s);
// run_median (lib.rs:1340-1372); flag: kMedianScratchWords zero-initialised words; flag[0] is set when a NaN is met (the
// reference panics), the rest is per-SM arrival counters of the one-pass kernel's staggered first wave
constexpr size_t kMedianScratchWords = 1 + 256;
cudaError_t median_launch(const float *cube, int64_t n2, int64_t start, int64_t end, float *out, unsigned *flag, cudaStream_t s);

// ---- FITS data unit -> f32 (ingest.cu) ----------------------------------------------------------
enum IngestFormat : int { // values match afk_fits_format (include/aflak_b200.h)
    ING_F32 = 0, // BITPIX -32
    ING_F64 = 1, // BITPIX -64
    ING_I16 = 2, // BITPIX 16
    ING_I32 = 3, // BITPIX 32
    ING_U32 = 4, // BITPIX 32 read as unsigned (fitrs FitsData::IntegersU32)
};
struct IngestArgs {
    const void *raw; // big-endian elements as they sit in the file; may alias out for 4-byte formats
    int fmt;
    int64_t n;       // elements
    int has_blank;
    int64_t blank;   // BLANK: integer formats only
    float bscale, bzero;
    float *out;
};
int ingest_element_bytes(int fmt); // 0 for an unknown format
cudaError_t ingest_launch(const IngestArgs &a, int sm_count, cudaStream_t s);

// ---- elementwise (K3) --------------------------------------------------------
enum MapOp : int {
    MAP_LINCOMP = 0,    // out = fl(fl(u*a) + fl(v*b))                     lib.rs:1288-1297
    MAP_RATIO = 1,      // out = u / v                                      unit.rs:863-880
    MAP_MULTIPLIER = 2, // out = powf(u,a) * powf(v,b)                      lib.rs:1619-1638
    MAP_EW = 3,         // out = EW(off=u, on=v, fl=a, max=b, sign=c)       lib.rs:1584-1596
    MAP_LOGSCALE = 4,   // out = ln(a*((u-b)/(c-b))+1)/ln(a)                lib.rs:1598-1611
    MAP_LOG10 = 5,      // out = log10(u)                                   lib.rs:1613-1617
    MAP_NEGATION = 6,   // out = -u                                         lib.rs:1640-1645
    MAP_NORMALIZE = 7,  // out = (u - a) / (b - a)                          lib.rs:269-273
    MAP_CLIP = 8,       // NaN where (ceil && u >= a) || (floor && u <= b)  lib.rs:1258-1274
    MAP_REPLACE_NAN = 9,// u is NaN ? a : u                                 lib.rs:1276-1286
    MAP_EMISSION_LINE = 10, // like EW without the division                 lib.rs run_create_emission_line_map
    MAP_VELOCITY = 11,  // out = ((u - a) * 3e5) / a                        lib.rs:509-526
    MAP_HISTTRANS = 12, // midtone transfer between min=a, max=b with parameters (c, d, e)   lib.rs:562-591
};

struct MapArgs {
    int op;
    const float *u;
    const float *v; // nullptr for unary ops
    float a, b, c;
    int flag0, flag1;
    int64_t n;
    float *out;
    float d = 0.0f, e = 0.0f; // MAP_HISTTRANS only
    const unsigned *bc_minmax = nullptr; // MAP_LOGSCALE only: minmax_kernel's device result {~key(min), key(max), any} replaces b, c
};
cudaError_t map_launch(const MapArgs &a, int sm_count, cudaStream_t s);

// apply_tone_curve (lib.rs:1858-1871). table_len <= kMaxToneTable. *oob_flag is set to 1 when an index
// exceeds the table (the reference panics there); the store is clamped so memory stays safe.
constexpr int kMaxToneTable = 4096;
cudaError_t tone_curve_launch(const float *img, int64_t n, const float *table_dev, int table_len, float *out,
                              unsigned *oob_flag, int sm_count, cudaStream_t s);

// per-pixel operators on 3-channel [3][n] images
enum Color3Op : int {
    C3_ARCSINH = 0, // apply_arcsinh_stretch (lib.rs:1873-1892): s = asinh(l*beta) / (l * asinh(beta)) in f64, l = (r+g+b)/3
    C3_TO_HSV = 1,  // color_image_to_hsv   (lib.rs:1795-1826)
};
cudaError_t color3_launch(int op, const float *img, int64_t n_per_channel, float beta, float *out, int sm_count, cudaStream_t s);

// ---- spatial reductions / gathers (spatial.cu) ---------------------------------------------------
// extract_wave: out[k] = sum over the region (offsets_dev: linear pixel offsets in list order, or nullptr = whole frame)
size_t wave_partial_bytes(int64_t nz, int64_t count, int *chunks_out, int64_t *chunk_out);
cudaError_t wave_launch(const float *cube, int64_t nz, int64_t n2, const int64_t *offsets_dev, int64_t count, int mode, void *partials,
                        float *out, cudaStream_t s);
cudaError_t extrude_launch(const float *cube, int64_t nz, int64_t n2, const int64_t *offsets_dev, int64_t count, float *out, int sm_count,
                           cudaStream_t s);
// *first_bad (device, initialised to ~0ull) receives the smallest map index that falls outside the cube
cudaError_t slice3d_launch(const float *cube, int64_t d0, int64_t d1, int64_t d2, const float *map_dev, int64_t n, float *out,
                           unsigned long long *first_bad, int sm_count, cudaStream_t s);

// ---- min/max (K4) -------------------------------------------------------------
// result[0] = ~key(min), result[1] = key(max), result[2] = number-of-non-NaN > 0 ; all via atomicMax, zero-init.
cudaError_t minmax_launch(const float *img, int64_t n, unsigned *result, int sm_count, cudaStream_t s);
// the same with the read-back folded in: see minmax_kernel (kernels.cu)
cudaError_t minmax_launch_readback(const float *img, int64_t n, unsigned *words4, unsigned *host_out, int sm_count, cudaStream_t s);
float minmax_decode_min(unsigned inv_key);
float minmax_decode_max(unsigned key);

// ---- display statistics (stats.cu) ----------------------------------------------------------------
// hist::histogram / histogram_color (aflak_plot/src/imshow/hist.rs:10-66): counts_dev is [nbins][channels] u64, zeroed
// ticket_dev / host_out: both null, or a zeroed device word and mapped pinned host memory for the counts -- the kernel's last
// CTA then stores the counts there and returns counts_dev and the ticket to zero (no memset before the next launch, no copy)
cudaError_t histogram_launch(const float *img, int64_t n_per_channel, int channels, float mn, float mx, int nbins,
                             unsigned long long *counts_dev, int sm_count, cudaStream_t s, unsigned *ticket_dev = nullptr,
                             unsigned long long *host_out = nullptr);
struct StatParams {
    float vmin, vmax; // d = (v - vmin) / (vmax - vmin)
    int use_vmed;     // select over |d - vmed| with vmed read from the previous statistic's state (lims.rs:87)
    int raw;          // select over the pixel values themselves and normalise the answer (see stat_value); never with use_vmed
    const unsigned *mm; // one-pass select only, may be null: device words {~key(min), key(max), any} that replace vmin / vmax / raw
};
struct StatState {    // device-resident state of one radix select; zero-initialised
    unsigned prefix, mask;
    unsigned long long kth, n, eq_count;
    unsigned above_min;   // smallest key above the 16-bit bucket of the answer; initialise to 0xFFFFFFFF
    unsigned cand_count;  // keys copied to the candidate buffer (> capacity: the buffer is incomplete)
    unsigned ticket;      // CTAs that finished the current pass
    int even, error;
    float result;
};
cudaError_t stat_select_launch(const float *img, int64_t n, const StatParams &p, const StatState *prev, StatState *state, unsigned *hist,
                               unsigned *cand, unsigned cand_cap, int sm_count, cudaStream_t s);
struct StatBracket {  // device-resident state of the one-pass (sample-bracketed) select; zero-initialised
    unsigned lo, hi;      // inclusive key bracket from the sample
    int shift;            // bin = (key - lo) >> shift, < 2048
    unsigned above_min;   // smallest kept key above hi; initialise to 0xFFFFFFFF
    unsigned next_min;    // smallest candidate in a bin above the chosen one; initialise to 0xFFFFFFFF
    unsigned cand_count, bin_cand_count, bin;
    unsigned ticket, ticket2;
    int even, error, fallback;
    unsigned pad;
    unsigned long long n, below, in_count, rank_in_bin, bin_count;
};
constexpr unsigned kStatBracketBins = 2048, kStatBracketBinBuf = 8192;
constexpr int64_t kStatBracketMinPixels = (int64_t)1 << 20; // below, the three radix passes are as fast (launch-bound)
cudaError_t stat_bracket_launch(const float *img, int64_t n, const StatParams &p, const StatState *prev, StatState *state, StatBracket *br,
                                unsigned *hist, unsigned *cand, unsigned cand_cap, unsigned *binbuf, int sm_count, cudaStream_t s);
cudaError_t zsample_launch(const float *img, int64_t ny, int64_t nx, int64_t stride, int64_t rows_s, int64_t cols_s, float *out,
                           cudaStream_t s);

// ---- gather (K6) ----------------------------------------------------------------
cudaError_t downsample_launch(const float *img, int64_t c, int64_t ny, int64_t nx, int64_t n, float *out, int64_t oy,
                              int64_t ox, cudaStream_t s);

// ---- colour mapping (K5) -------------------------------------------------------
struct ColormapArgs {
    const float *img;     // [ny][nx], or [3][ny][nx] when rgb3
    int64_t ny, nx;
    const uint32_t *lut;  // 65536 entries (or 3 x 65536 when rgb3), packed r | g<<8 | b<<16 | 0xFF<<24
    float vmin[3], vmax[3];
    float lims[3][3];     // (min, mid, max) per channel table
    int rgb3;             // make_raw_image_rgb (image.rs:51-84)
    int flip_y;           // coerce_to_array_view2 (image.rs:123-129)
    int channels;         // 3 = packed RGB8 (reference ClientFormat::U8U8U8), 4 = RGBA8
    // optional fused upstream chain (BASELINE C4): logscale -> normalize -> tone curve, applied per pixel before the LUT
    int chain;            // 0 none, 1 logscale(a, lmin, lmax) -> normalize(nmin, nmax) -> tone(table)
    float ls_a, ls_min, ls_max, nm_min, nm_max;
    const float *tone_table;
    int tone_len;
    unsigned *oob_flag;
    const unsigned *ls_minmax; // chain only, may be null: minmax_kernel's result {~key(min), key(max), any} replaces ls_min / ls_max
    uint8_t *out;         // [ny][nx][channels]
};
cudaError_t colormap_launch(const ColormapArgs &a, int sm_count, cudaStream_t s);

// ---- synthetic generator (same arithmetic as oracle/aflak_oracle.c orc_gen_voxel) ----
cudaError_t gen_cube_launch(uint64_t seed, int64_t nz, int64_t ny, int64_t nx, int64_t y0, int64_t rows, int pathology,
                            float *out, cudaStream_t s);

// display kernels: fast forms (default) or the reference's IEEE sequence for every pixel
bool display_exact();
void set_display_exact(bool on);

// number of kernel launches issued by this library since load (bench.py's gpu_launches)
uint64_t launch_count();
void count_launch(); // for the other translation units

} // namespace afk
