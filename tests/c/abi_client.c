/* abi_client.c -- a plain C99 client of include/aflak_b200.h, written the way the Rust `-sys` binding of INTEGRATION.md
 * would use it: no Python, no C++, only the exported entry points.
 *
 *   gcc -std=c99 -O2 -I include tests/c/abi_client.c -L aflak_b200 -laflak_b200 -Wl,-rpath,$PWD/aflak_b200 -lm -o abi_client
 *   ./abi_client            (needs a CUDA device; exits 77 when there is none)
 *
 * It (1) walks the registry like `TRANSFORMATIONS.iter()` (aflak_primitives/src/lib.rs:141-758),
 *    (2) calls `integral` through the PlainFunction pointer of its descriptor (aflak_cake/src/transform.rs:78) on a cube
 *        whose answer is known in closed form (cube[k,y,x] = k: sum over [a,b) = (a+b-1)(b-a)/2),
 *    (3) checks the reference's error wording for a bad range (precond.rs:3-15),
 *    (4) evaluates the equivalent-width template tail through the graph evaluator with a cached device cube
 *        (aflak_cake/src/dst/compute.rs:153-274) and re-evaluates it after a constant changes,
 *    (5) prints the host-side cost per call of the small-cube cases (BASELINE configs 0 and 1) measured from C.
 * Test infrastructure: tests/test_abi_cpu.py compiles it, tests/test_gpu_graph.py runs it.
 */
#define _POSIX_C_SOURCE 200809L
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#include "aflak_b200.h"

static int failures = 0;
#define CHECK(cond, ...)                                                                                                         \
    do {                                                                                                                         \
        if (!(cond)) {                                                                                                           \
            fprintf(stderr, "FAIL %s:%d: ", __FILE__, __LINE__);                                                                 \
            fprintf(stderr, __VA_ARGS__);                                                                                        \
            fprintf(stderr, "\n");                                                                                               \
            ++failures;                                                                                                          \
        }                                                                                                                        \
    } while (0)

static double now_us(void) {
    struct timespec t;
    clock_gettime(CLOCK_MONOTONIC, &t);
    return t.tv_sec * 1e6 + t.tv_nsec * 1e-3;
}

/* integral(image, start, end) through the descriptor's function pointer; returns the 2-D image value (owned) or NULL */
static afk_value *call_integral(const afk_transform_desc *t, afk_value *cube, int64_t start, int64_t end, afk_result out[3]) {
    afk_value *a = afk_value_integer(start), *b = afk_value_integer(end);
    const afk_value *in[3];
    in[0] = cube; in[1] = a; in[2] = b;
    memset(out, 0, 3 * sizeof out[0]);
    t->f(in, 3, out, 3);
    afk_value_release(a);
    afk_value_release(b);
    return out[0].status == AFK_OK ? out[0].value : NULL;
}

int main(void) {
    if (afk_device_count() < 1) {
        fprintf(stderr, "abi_client: no CUDA device (the library has no CPU fallback)\n");
        return 77;
    }
    CHECK(afk_abi_version() == 1, "abi version %d", afk_abi_version());
    CHECK(afk_set_device(0) == AFK_OK, "set_device: %s", afk_last_error());

    /* (1) registry walk */
    size_t n_t = afk_transform_count(), with_image_input = 0;
    for (size_t i = 0; i < n_t; ++i) {
        const afk_transform_desc *t = afk_transform_at(i);
        CHECK(t && t->name && t->f && afk_get_transform(t->name) == t, "descriptor %zu", i);
        for (size_t k = 0; t && k < t->n_inputs; ++k)
            if (t->inputs[k].type_id == AFK_T_IMAGE) { ++with_image_input; break; }
    }
    CHECK(n_t >= 40 && with_image_input >= 30, "registry: %zu transforms, %zu with an Image input", n_t, with_image_input);
    const afk_transform_desc *integral = afk_get_transform("integral");
    CHECK(integral && integral->n_inputs == 3 && integral->n_outputs == 3 && integral->major == 1, "integral descriptor");
    if (!integral) return 1;

    /* (2) closed-form known answer: cube[k,y,x] = k */
    enum { NZ = 300, NY = 37, NX = 52 };
    float *host = (float *)malloc(sizeof(float) * NZ * NY * NX);
    for (int k = 0; k < NZ; ++k)
        for (int p = 0; p < NY * NX; ++p) host[(size_t)k * NY * NX + p] = (float)k;
    const int64_t shape[3] = {NZ, NY, NX};
    afk_image *cube_img = afk_image_upload(host, 3, shape);
    CHECK(cube_img != NULL, "upload: %s", afk_last_error());
    if (!cube_img) return 1;
    afk_value *cube = afk_value_image(cube_img);
    afk_image_release(cube_img); /* the value holds its own reference, like Arc<IOValue> in the cake cache */
    afk_result out[3];
    const int64_t a = 17, b = 211;
    afk_value *img = call_integral(integral, cube, a, b, out);
    CHECK(img != NULL, "integral: %s", out[0].message ? out[0].message : "?");
    if (img) {
        afk_image *r = afk_value_get_image(img);
        CHECK(afk_image_ndim(r) == 2 && afk_image_shape(r)[0] == NY && afk_image_shape(r)[1] == NX, "integral shape");
        float *res = (float *)malloc(sizeof(float) * NY * NX);
        CHECK(afk_image_download(r, res) == AFK_OK, "download: %s", afk_last_error());
        const float want = (float)((a + b - 1) * (b - a) / 2);
        int bad = 0;
        for (int p = 0; p < NY * NX; ++p) bad += res[p] != want;
        CHECK(bad == 0, "integral of k over [%lld,%lld): %d pixels differ from %g (got %g)", (long long)a, (long long)b, bad, want, res[0]);
        float middle = 0, width = 0;
        CHECK(afk_value_get_float(out[1].value, &middle) == AFK_OK && afk_value_get_float(out[2].value, &width) == AFK_OK, "scalar outputs");
        CHECK(middle == ((float)a + (float)b) / 2.0f && width == (float)(b - a), "middle %g width %g", middle, width);
        free(res);
    }
    for (int i = 0; i < 3; ++i) afk_result_clear(&out[i]);

    /* (3) the reference's error wording (precond.rs:3-15) comes back as Err(IOErr::UnexpectedInput), never as an abort */
    CHECK(call_integral(integral, cube, -1, 5, out) == NULL && out[0].status == AFK_ERR_UNEXPECTED_INPUT, "negative start accepted");
    CHECK(out[0].message && strstr(out[0].message, "'start' must be positive, but got -1"), "message: %s", out[0].message ? out[0].message : "(null)");
    for (int i = 0; i < 3; ++i) afk_result_clear(&out[i]);
    CHECK(call_integral(integral, cube, 9, 9, out) == NULL && out[0].message && strstr(out[0].message, "Expected start < end, but got 9 < 9"),
          "message: %s", out[0].message ? out[0].message : "(null)");
    for (int i = 0; i < 3; ++i) afk_result_clear(&out[i]);

    /* (4) evaluator with a device-handle cache: average -> image_min_max -> convert_to_logscale on the cached cube */
    afk_graph *g = afk_graph_new();
    afk_value *v_s = afk_value_integer(10), *v_e = afk_value_integer(40), *v_a = afk_value_float(1000.0f);
    const int64_t c_cube = afk_graph_add_constant(g, cube), c_s = afk_graph_add_constant(g, v_s), c_e = afk_graph_add_constant(g, v_e),
                  c_a = afk_graph_add_constant(g, v_a);
    const int64_t t_avg = afk_graph_add_transform(g, "average"), t_mm = afk_graph_add_transform(g, "image_min_max"),
                  t_ls = afk_graph_add_transform(g, "convert_to_logscale");
    CHECK(t_avg >= 0 && t_mm >= 0 && t_ls >= 0, "add_transform");
    CHECK(afk_graph_connect(g, c_cube, 0, t_avg, 0) == AFK_OK && afk_graph_connect(g, c_s, 0, t_avg, 1) == AFK_OK &&
              afk_graph_connect(g, c_e, 0, t_avg, 2) == AFK_OK,
          "connect average: %s", afk_last_error());
    CHECK(afk_graph_connect(g, t_avg, 0, t_mm, 0) == AFK_OK && afk_graph_connect(g, t_avg, 0, t_ls, 0) == AFK_OK &&
              afk_graph_connect(g, c_a, 0, t_ls, 1) == AFK_OK && afk_graph_connect(g, t_mm, 0, t_ls, 2) == AFK_OK &&
              afk_graph_connect(g, t_mm, 1, t_ls, 3) == AFK_OK,
          "connect logscale: %s", afk_last_error());
    /* an Image output cannot feed an Integer slot (dst/build.rs:45-88) */
    CHECK(afk_graph_connect(g, t_avg, 0, t_avg, 1) != AFK_OK, "ill-typed edge accepted");
    afk_result r;
    memset(&r, 0, sizeof r);
    CHECK(afk_graph_compute(g, t_ls, 0, &r) == AFK_OK && r.value && afk_value_type(r.value) == AFK_T_IMAGE, "compute: %s", r.message ? r.message : afk_last_error());
    if (r.value) {
        /* average of k over [10,40) = 24.5 everywhere: min == max, so (v - min) / (max - min) = 0/0 = NaN for every pixel */
        float px[NY * NX];
        CHECK(afk_image_download(afk_value_get_image(r.value), px) == AFK_OK, "download");
        CHECK(px[0] != px[0] && px[NY * NX - 1] != px[NY * NX - 1], "zero-range logscale must be NaN like the reference (got %g)", px[0]);
    }
    afk_result_clear(&r);
    const uint64_t calls0 = afk_graph_calls(g);
    CHECK(afk_graph_compute(g, t_ls, 0, &r) == AFK_OK, "recompute");
    afk_result_clear(&r);
    CHECK(afk_graph_calls(g) == calls0, "a cached output was recomputed (%llu -> %llu calls)", (unsigned long long)calls0, (unsigned long long)afk_graph_calls(g));
    afk_value *v_e2 = afk_value_integer(41);
    CHECK(afk_graph_set_constant(g, c_e, v_e2) == AFK_OK, "set_constant");
    CHECK(afk_graph_compute(g, t_ls, 0, &r) == AFK_OK && afk_graph_calls(g) > calls0, "a changed constant did not invalidate its descendants");
    afk_result_clear(&r);

    /* (5) host-side cost per call from C (the Python wrapper used by bench.py adds ~10 us of its own) */
    {
        const int reps = 2000;
        afk_synchronize();
        double t0 = now_us();
        for (int i = 0; i < reps; ++i) {
            afk_value *im = call_integral(integral, cube, 100, 120, out);
            (void)im;
            for (int k = 0; k < 3; ++k) afk_result_clear(&out[k]);
        }
        afk_synchronize();
        const double per_call = (now_us() - t0) / reps;
        t0 = now_us();
        for (int i = 0; i < reps; ++i) {
            afk_value *ve = afk_value_integer(40 + (i & 1));
            afk_graph_set_constant(g, c_e, ve);
            afk_value_release(ve);
            afk_graph_compute(g, t_ls, 0, &r);
            afk_result_clear(&r);
        }
        afk_synchronize();
        const double per_eval = (now_us() - t0) / reps;
        printf("{\"c_abi_client\": {\"cube\": [%d, %d, %d], \"integral_20_frames_us_per_call\": %.2f, "
               "\"graph_reevaluation_average_minmax_logscale_us\": %.2f, \"kernel_launches\": %llu}}\n",
               NZ, NY, NX, per_call, per_eval, (unsigned long long)afk_launch_count());
    }

    afk_value_release(v_e2);
    afk_value_release(v_s);
    afk_value_release(v_e);
    afk_value_release(v_a);
    afk_graph_free(g);
    afk_value_release(cube);
    free(host);
    if (failures) {
        fprintf(stderr, "abi_client: %d check(s) failed\n", failures);
        return 1;
    }
    printf("abi_client ok\n");
    return 0;
}
