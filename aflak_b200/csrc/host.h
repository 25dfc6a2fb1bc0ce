// host.h -- internal declarations of the C++ host layer behind include/aflak_b200.h.
#pragma once
#include <cuda_runtime.h>

#include <atomic>
#include <initializer_list>
#include <map>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/aflak_b200.h"
#include "kernels.h"

namespace afk {

// ---- errors ----------------------------------------------------------------------------------------
struct Error {
    afk_status status = AFK_OK;
    std::string message;
    explicit operator bool() const { return status != AFK_OK; }
};
inline Error ok() { return Error{}; }
Error make_error(afk_status st, const std::string &msg);
Error cuda_error(cudaError_t e, const char *what);
std::string fmt(const char *f, ...) __attribute__((format(printf, 1, 2)));
void set_last_error(const std::string &msg);
afk_status fail(const Error &e); // records the message for afk_last_error(), returns the status

#define AFK_CUDA(call)                                        \
    do {                                                      \
        cudaError_t e__ = (call);                             \
        if (e__ != cudaSuccess) return cuda_error(e__, #call); \
    } while (0)

// ---- per-thread execution context -----------------------------------------------------------------
struct StreamScratch {
    ReduceWorkspace ws{nullptr, 0, nullptr, 0};
    // display chain: the last tone table uploaded on this stream (re-used while the host passes the same entries) and the
    // three words image_min_max leaves on the device for the chain kernel.  Stream ordered, owned by the calling thread.
    float *tone_dev = nullptr;
    size_t tone_cap = 0;
    std::vector<float> tone_host;
    unsigned *minmax_words = nullptr;
    unsigned *minmax_self = nullptr; // op_min_max: four self-resetting words (minmax_launch_readback), zeroed once
    unsigned long long *hist_self = nullptr; // histogram_run: kHistSelfEntries self-resetting counts + the ticket word, zeroed once
};
constexpr size_t kHistSelfEntries = 768; // 256 bins x 3 channels (hist.rs:36-66)

struct ThreadCtx {
    int device = -1;
    cudaStream_t own[16] = {};
    cudaStream_t external = nullptr;
    std::map<std::pair<int, cudaStream_t>, StreamScratch> scratch;
    int sum_mode = SUM_INHERIT; // per-thread override of the process default (afk_set_default_sum_mode)
    bool profile = false;
    std::vector<std::pair<cudaEvent_t, cudaEvent_t>> prof_events;
    uint64_t prof_launches = 0;
    std::string last_error;
    // Deferred reference panics: a word of mapped pinned host memory per device that kernels on this thread's stream set
    // where the reference would panic (tone-curve index beyond the table).  Read after any host synchronisation.
    unsigned *panic_word[16] = {};
    size_t panic_tone_len = 0; // table length of the last chain / tone-curve launch, for the message
    // One page of mapped pinned host memory per thread for the few words an operator reads back (limits, order statistics):
    // an asynchronous copy into it (or a kernel's own store) plus ONE stream synchronisation, instead of one synchronising
    // pageable copy per value.  Only valid between the enqueue and the synchronisation of one operator call on this thread.
    char *readback = nullptr;
};
constexpr size_t kReadbackBytes = 8192; // [0, 512) order statistics | [512, 1024) min/max | [1024, 8192) histogram counts, zscale samples
Error readback_page(char **out);
ThreadCtx &tls();
int effective_sum_mode(); // the thread's override, else the process-wide default
// make sure a device is current and return the stream operators on this thread use
Error current_stream(cudaStream_t *out, int *device);
int sm_count(int device);
// device-usable pointer (UVA) to this thread's panic word for `device`; the kernel stores 1 to it with a plain store
Error panic_word_get(int device, unsigned **out);
// after the stream has been synchronised: the pending panic of this thread on `device` as an error (and clear it), or ok
Error panic_word_take(int device);
cudaEvent_t event_get();
void event_put(cudaEvent_t e);

// ---- images ----------------------------------------------------------------------------------------
struct AxisMeta {
    bool has_name = false;
    std::string name;
    std::string unit; // "" == Unit::None
    bool operator==(const AxisMeta &o) const { return has_name == o.has_name && name == o.name && unit == o.unit; }
};

struct ImageMeta { // host-side part of WcsArray (unit.rs:190-242); the WCS transform itself is an opaque blob
    bool has_unit = false;
    std::string unit;
    bool homogeneous = true;
    bool has_tag = false;
    std::string tag;
    bool has_meta = false;
    AxisMeta axes[4];
    std::vector<double> wcs; // opaque: CRVAL/CRPIX/CDELT per kept axis, copied or trimmed by make_slice
    bool meta_equal(const ImageMeta &o) const;
};

} // namespace afk

struct afk_image {
    std::atomic<int> rc{1};
    float *ptr = nullptr;
    size_t bytes = 0;
    bool owned = true;
    bool ipc_peer = false;
    bool track_readers = false; // borrowed memory whose owner (afk_gather) orders its reuse after recorded readers
    afk_image *parent = nullptr; // a view keeps its parent alive
    int device = 0;
    int ndim = 0;
    int64_t shape[AFK_MAX_NDIM] = {0, 0, 0, 0};
    cudaStream_t stream = nullptr; // producer stream
    cudaEvent_t ready = nullptr;   // recorded after the producing kernel
    std::mutex mu;
    std::vector<cudaEvent_t> readers; // last reads from other streams, waited on before the memory is freed
    afk::ImageMeta meta;
    // Range memo: four words {~key(min), key(max), any non-NaN, 0} kept on the device right after the pixels of an image
    // that image_new allocated.  The kernel that writes the pixels may fill them (range_valid); image_min_max, get_vmin /
    // get_vmax and the fused consumers then read the range there instead of making a pass over the image.  Images are
    // immutable values once ready, so a valid memo stays valid.
    unsigned *range = nullptr;
    bool range_valid = false;

    int64_t len() const {
        int64_t n = 1;
        for (int i = 0; i < ndim; ++i) n *= shape[i];
        return n;
    }
    int64_t frame_len() const { // product of all but the first axis
        int64_t n = 1;
        for (int i = 1; i < ndim; ++i) n *= shape[i];
        return n;
    }
};

struct afk_value {
    std::atomic<int> rc{1};
    afk_type type = AFK_T_INTEGER;
    int64_t i = 0;
    float f[9] = {0};
    bool b = false;
    std::string s;
    afk_image *img = nullptr;
    std::vector<float> arr; // ToneCurve arr
    std::vector<float> cp;  // ToneCurve control points (x,y pairs)
    std::vector<std::string> paths; // PATHS::FileList
    afk_fits *fits = nullptr;       // Arc<fitrs::Fits>
    bool roi_all = true;            // ROI::All, else ROI::PixelList(roi_xy)
    std::vector<uint64_t> roi_xy;   // (i, j) = (x, y) pairs (roi.rs:5-10)
    std::vector<float> map;         // Map2dTo3dCoords: Array2<[f32; 3]>, [map_rows][map_cols][3]
    int64_t map_rows = 0, map_cols = 0;
    // lazily built by slice_3d_to_2d and kept with the (immutable) value: device copy of the mesh and its axis analysis
    std::mutex map_mu;
    float *map_dev = nullptr;
    int map_dev_device = -1;
    int map_sliced = -1; // -1 not analysed, 0 None, 1 Some(map_axis, map_start, map_step)
    int map_axis[2] = {0, 0};
    float map_start[2] = {0, 0}, map_step[2] = {0, 0};    // lazily built by extract_wave / extrude: device copy of ROI::filter's offsets for a [roi_ny, roi_nx] frame (same mutex)
    int64_t *roi_dev = nullptr;
    int roi_dev_device = -1;
    int64_t roi_ny = -1, roi_nx = -1, roi_count = 0;
};

namespace afk {

// allocate an output image on the calling thread's stream
Error image_new(int ndim, const int64_t *shape, afk_image **out);
// order `stream` after the producer of img (no-op on the same stream)
Error image_wait(const afk_image *img, cudaStream_t stream);
// note that `stream` has enqueued a read of img
void image_note_read(const afk_image *img, cudaStream_t stream);
// mark img as produced on stream
Error image_mark_ready(afk_image *img, cudaStream_t stream);

// get (growing if needed) the reduce workspace of (device, stream)
Error reduce_workspace(const ReducePlan &plan, int device, cudaStream_t stream, ReduceWorkspace *ws);

// small zero-initialised device scratch, stream ordered
Error scratch_u32(cudaStream_t s, size_t n, unsigned **out);
void scratch_free(cudaStream_t s, void *p);

// ---- operators on handles (ops.cpp) -----------------------------------------------------------------
Error op_reduce(const afk_image *cube, int nwin, const int64_t *starts, const int64_t *ends, int epilogue,
                const float *params, float ew_sign, bool want_minmax, afk_image **out, float *minmax);
Error op_map(int op, const afk_image *u, const afk_image *v, float a, float b, float c, int flag0, int flag1, afk_image **out);
Error op_min_max(const afk_image *img, float *mn, float *mx, bool *any);

// unit algebra (unit.rs:39-163, 454-495)
std::string unit_mul(bool has_a, const std::string &a, bool has_b, const std::string &b, bool *has_out);
std::string unit_div(bool has_a, const std::string &a, bool has_b, const std::string &b, bool *has_out);

// WcsArray::pix2world (unit.rs:335-342) on the linear per-axis WCS carried in the meta blob
float image_pix2world(const afk_image *img, int axis, float pixel);

// result helpers shared by the operator translation units (ops.cpp, fits.cpp)
void set_ok(afk_result *r, afk_value *v);
void set_err(afk_result *r, const Error &e);
void set_image(afk_result *r, afk_image *img); // takes over the caller's reference
bool args_match(const afk_value *const *in, size_t n_in, std::initializer_list<afk_type> types, afk_result *out, size_t n_out);
Error unexpected(const std::string &m);
Error try_into_unsigned(const char *name, int64_t v);

// operators of ops_spatial.cpp
#define AFK_DECL_OP(name) afk_status name(const afk_value *const *in, size_t n_in, afk_result *out, size_t n_out)
AFK_DECL_OP(op_make_float3);
AFK_DECL_OP(op_make_plane3d);
AFK_DECL_OP(op_slice_3d_to_2d);
AFK_DECL_OP(op_extrude);
AFK_DECL_OP(op_extract_wave);
AFK_DECL_OP(op_create_velocity_field_map);
AFK_DECL_OP(op_histogram_transformation);
AFK_DECL_OP(op_histogram_transformation_rgb);
AFK_DECL_OP(op_apply_arcsinh_stretch);
AFK_DECL_OP(op_color_image_to_hsv);
AFK_DECL_OP(op_color_image_from_rgb);
Error are_same_dim(const char *n1, const afk_image *a, const char *n2, const afk_image *b);
Error same_device(const afk_image *a, int device);
std::string rust_f32_display(float f);

// FITS operators (fits.cpp)
afk_status op_open_fits(const afk_value *const *in, size_t n_in, afk_result *out, size_t n_out);
afk_status op_fits_to_image(const afk_value *const *in, size_t n_in, afk_result *out, size_t n_out);

void registry_init();

} // namespace afk
