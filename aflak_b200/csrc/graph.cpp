// graph.cpp -- minimal DAG evaluator with a device-handle cache.
//
// Restates the *contract* of aflak_cake's DST::compute + Cache (aflak_cake/src/dst/compute.rs:153-274,
// cache.rs:18-148, dst/mod.rs:65-73, dst/build.rs:45-88), not its implementation: nodes are registry transforms or
// constants, a node's outputs are memoised with the logical time of the computation and reused iff that time is
// not older than `updated_on` of the node or any ancestor.  What is cached are refcounted afk_value's whose Image
// payloads are device handles -- never host copies (north_star subsystem 3).
// It also recognises the two sub-graphs BASELINE names and substitutes the fused kernels for them.
#include <cstring>
#include <functional>
#include <set>

#include "host.h"

namespace afk {

struct Edge {
    int64_t from_t = -1;
    size_t out_i = 0;
};

struct CacheBox { // cache.rs:79-82
    bool valid = false;
    uint64_t time = 0;
    std::vector<afk_result> values; // one Result per output, values retained by the cache
};

struct Node {
    bool is_constant = false;
    const afk_transform_desc *t = nullptr;
    afk_value *constant = nullptr;
    uint64_t updated_on = 0;
    std::vector<Edge> inputs; // per input slot; from_t < 0 = unattached
    CacheBox cache;
};

static std::atomic<uint64_t> g_clock{1};
static uint64_t now() { return g_clock.fetch_add(1); }

static void copy_result(const afk_result &src, afk_result *dst) {
    dst->status = src.status;
    dst->value = src.value;
    if (dst->value) afk_value_retain(dst->value);
    dst->message = src.message ? strdup(src.message) : nullptr;
}

} // namespace afk

struct afk_graph {
    std::vector<afk::Node *> nodes;
    std::mutex mu;
    uint64_t calls = 0;
    bool fusion = true;
};

namespace afk {

static size_t n_outputs(const Node *n) { return n->is_constant ? 1 : n->t->n_outputs; }
static afk_type output_type(const Node *n, size_t i) { return n->is_constant ? n->constant->type : n->t->outputs[i]; }

static Node *node_at(afk_graph *g, int64_t t) { return (t >= 0 && (size_t)t < g->nodes.size()) ? g->nodes[(size_t)t] : nullptr; }

// dst/mod.rs:65-73: max(updated_on) over the node and all its ancestors
static uint64_t updated_on(afk_graph *g, int64_t t, std::set<int64_t> &seen) {
    Node *n = node_at(g, t);
    if (!n || !seen.insert(t).second) return 0;
    uint64_t u = n->updated_on;
    for (const Edge &e : n->inputs)
        if (e.from_t >= 0) u = std::max(u, updated_on(g, e.from_t, seen));
    return u;
}

static bool depends_on(afk_graph *g, int64_t t, int64_t target) {
    if (t == target) return true;
    Node *n = node_at(g, t);
    if (!n) return false;
    for (const Edge &e : n->inputs)
        if (e.from_t >= 0 && depends_on(g, e.from_t, target)) return true;
    return false;
}

static void clear_cache(Node *n) {
    for (afk_result &r : n->cache.values) afk_result_clear(&r);
    n->cache.values.clear();
    n->cache.valid = false;
}

static void fill_error(std::vector<afk_result> &res, afk_status st, const std::string &msg) {
    for (afk_result &r : res) {
        afk_result_clear(&r);
        r.status = st;
        r.message = strdup(msg.c_str());
    }
}

// ---- fusion: pattern-match the sub-graphs BASELINE names --------------------------------------------------------------
struct AvgNode {
    int64_t cube_t;
    size_t cube_out;
    int64_t start, end;
};

static bool const_int(afk_graph *g, const Node *n, size_t slot, int64_t *out) {
    const Edge &e = n->inputs[slot];
    const afk_value *v = nullptr;
    if (e.from_t >= 0) {
        Node *src = node_at(g, e.from_t);
        if (!src || !src->is_constant) return false;
        v = src->constant;
    } else {
        v = n->t->inputs[slot].default_value;
    }
    if (!v) return false;
    if (v->type == AFK_T_INTEGER) { *out = v->i; return true; }
    return false;
}

static bool const_scalar(afk_graph *g, const Node *n, size_t slot, afk_type want, const afk_value **out) {
    const Edge &e = n->inputs[slot];
    const afk_value *v = nullptr;
    if (e.from_t >= 0) {
        Node *src = node_at(g, e.from_t);
        if (!src || !src->is_constant) return false;
        v = src->constant;
    } else {
        v = n->t->inputs[slot].default_value;
    }
    if (!v) return false;
    if (v->type != want && !(want == AFK_T_FLOAT && v->type == AFK_T_INTEGER)) return false;
    *out = v;
    return true;
}

static float scalar_as_float(const afk_value *v) { return v->type == AFK_T_INTEGER ? (float)v->i : v->f[0]; }

static bool match_average(afk_graph *g, const Edge &e, AvgNode *out) {
    if (e.from_t < 0 || e.out_i != 0) return false;
    Node *n = node_at(g, e.from_t);
    if (!n || n->is_constant || strcmp(n->t->name, "average") != 0) return false;
    if (n->inputs[0].from_t < 0) return false;
    out->cube_t = n->inputs[0].from_t;
    out->cube_out = n->inputs[0].out_i;
    return const_int(g, n, 1, &out->start) && const_int(g, n, 2, &out->end);
}

} // namespace afk

using namespace afk;

static afk_status compute_node(afk_graph *g, int64_t t, std::vector<afk_result> **out);

// Try the fused kernels.  Returns true when `res` has been filled.
static bool try_fused(afk_graph *g, Node *n, std::vector<afk_result> &res) {
    if (!g->fusion || n->is_constant) return false;
    const char *name = n->t->name;
    if (strcmp(name, "ratio") == 0) {
        // ratio(average(cube, s0, e0), average(cube, s1, e1))  -- BASELINE config 3
        AvgNode a, b;
        if (!match_average(g, n->inputs[0], &a) || !match_average(g, n->inputs[1], &b)) return false;
        if (a.cube_t != b.cube_t || a.cube_out != b.cube_out) return false;
        std::vector<afk_result> *cube_res = nullptr;
        if (compute_node(g, a.cube_t, &cube_res) != AFK_OK || (*cube_res)[a.cube_out].status != AFK_OK) return false;
        afk_value *cv = (*cube_res)[a.cube_out].value;
        if (!cv || cv->type != AFK_T_IMAGE) return false;
        afk_image *o = nullptr;
        if (afk_fused_average_ratio(cv->img, a.start, a.end, b.start, b.end, &o) != AFK_OK) return false; // unfused path reports the error
        res[0].status = AFK_OK;
        res[0].value = afk_value_image(o);
        afk_image_release(o);
        g->calls += 1;
        return true;
    }
    if (strcmp(name, "convert_to_logscale") == 0) {
        // convert_to_logscale(x, a, image_min_max(x).0, image_min_max(x).1): the range stays on the device
        const Edge &ex = n->inputs[0], &emin = n->inputs[2], &emax = n->inputs[3];
        if (ex.from_t < 0 || emin.from_t < 0 || emin.from_t != emax.from_t || emin.out_i != 0 || emax.out_i != 1) return false;
        Node *mmn = node_at(g, emin.from_t);
        if (!mmn || mmn->is_constant || strcmp(mmn->t->name, "image_min_max") != 0) return false;
        if (mmn->inputs[0].from_t != ex.from_t || mmn->inputs[0].out_i != ex.out_i) return false;
        const afk_value *av;
        if (!const_scalar(g, n, 1, AFK_T_FLOAT, &av)) return false;
        std::vector<afk_result> *src = nullptr;
        if (compute_node(g, ex.from_t, &src) != AFK_OK || (*src)[ex.out_i].status != AFK_OK) return false;
        afk_value *xv = (*src)[ex.out_i].value;
        if (!xv || xv->type != AFK_T_IMAGE) return false;
        afk_image *o = nullptr;
        if (afk_fused_logscale_minmax(xv->img, scalar_as_float(av), &o) != AFK_OK) return false;
        res[0].status = AFK_OK;
        res[0].value = afk_value_image(o);
        afk_image_release(o);
        g->calls += 1;
        return true;
    }
    if (strcmp(name, "create_equivalent_width") == 0) {
        // create_equivalent_width(linear_composition(avg(off1), avg(off2), r2, r1), avg(on), fl, max, emission) with
        // (r2, r1) = ratio_from_bands(mid_on, mid_off1, mid_off2)  -- src/templates.rs:143-463
        AvgNode on, off1, off2;
        if (!match_average(g, n->inputs[1], &on)) return false;
        const Edge &le = n->inputs[0];
        if (le.from_t < 0 || le.out_i != 0) return false;
        Node *lc = node_at(g, le.from_t);
        if (!lc || lc->is_constant || strcmp(lc->t->name, "linear_composition") != 0) return false;
        if (!match_average(g, lc->inputs[0], &off1) || !match_average(g, lc->inputs[1], &off2)) return false;
        if (on.cube_t != off1.cube_t || on.cube_t != off2.cube_t || on.cube_out != off1.cube_out || on.cube_out != off2.cube_out) return false;
        // a and b must come from one ratio_from_bands node fed by the three `middle` outputs
        const Edge &ea = lc->inputs[2], &eb = lc->inputs[3];
        if (ea.from_t < 0 || ea.from_t != eb.from_t || ea.out_i != 0 || eb.out_i != 1) return false;
        Node *rb = node_at(g, ea.from_t);
        if (!rb || rb->is_constant || strcmp(rb->t->name, "ratio_from_bands") != 0) return false;
        const int64_t want_src[3] = {n->inputs[1].from_t, lc->inputs[0].from_t, lc->inputs[1].from_t};
        for (int k = 0; k < 3; ++k)
            if (rb->inputs[k].from_t != want_src[k] || rb->inputs[k].out_i != 1) return false;
        const afk_value *fl, *mx, *em;
        if (!const_scalar(g, n, 2, AFK_T_FLOAT, &fl) || !const_scalar(g, n, 3, AFK_T_FLOAT, &mx) || !const_scalar(g, n, 4, AFK_T_BOOL, &em)) return false;
        std::vector<afk_result> *cube_res = nullptr;
        if (compute_node(g, on.cube_t, &cube_res) != AFK_OK || (*cube_res)[on.cube_out].status != AFK_OK) return false;
        afk_value *cv = (*cube_res)[on.cube_out].value;
        if (!cv || cv->type != AFK_T_IMAGE) return false;
        afk_image *o = nullptr;
        if (afk_fused_equivalent_width(cv->img, on.start, on.end, off1.start, off1.end, off2.start, off2.end, scalar_as_float(fl),
                                       scalar_as_float(mx), em->b ? 1 : 0, &o, nullptr) != AFK_OK)
            return false;
        res[0].status = AFK_OK;
        res[0].value = afk_value_image(o);
        afk_image_release(o);
        g->calls += 1;
        return true;
    }
    return false;
}

// DST::_compute (dst/compute.rs:179-274) with CacheRef::compute (cache.rs:121-147)
static afk_status compute_node(afk_graph *g, int64_t t, std::vector<afk_result> **out) {
    Node *n = node_at(g, t);
    if (!n) return fail(make_error(AFK_ERR_ARGUMENT, fmt("Node #%lld not found", (long long)t)));
    std::set<int64_t> seen;
    const uint64_t upd = updated_on(g, t, seen);
    if (n->cache.valid && n->cache.time >= upd) { // cache hit (cache.rs:130-136)
        *out = &n->cache.values;
        return AFK_OK;
    }
    std::vector<afk_result> res(n_outputs(n), afk_result{AFK_OK, nullptr, nullptr});
    if (n->is_constant) {
        res[0].value = n->constant;
        afk_value_retain(n->constant);
    } else if (!try_fused(g, n, res)) {
        const afk_transform_desc *td = n->t;
        std::vector<const afk_value *> args(td->n_inputs, nullptr);
        bool failed = false;
        for (size_t i = 0; i < td->n_inputs && !failed; ++i) {
            const Edge &e = n->inputs[i];
            if (e.from_t >= 0) {
                std::vector<afk_result> *pr = nullptr;
                afk_status st = compute_node(g, e.from_t, &pr);
                if (st != AFK_OK || e.out_i >= pr->size()) {
                    fill_error(res, AFK_ERR_ARGUMENT, fmt("%s\n    in node #%lld %s", afk_last_error(), (long long)t, td->name));
                    failed = true;
                } else if ((*pr)[e.out_i].status != AFK_OK) {
                    // ErrorStack (dst/compute.rs:240-247): every output of this node carries the parent's error
                    const afk_result &p = (*pr)[e.out_i];
                    fill_error(res, p.status, fmt("%s\n    in node #%lld %s", p.message ? p.message : "", (long long)t, td->name));
                    failed = true;
                } else {
                    args[i] = (*pr)[e.out_i].value;
                }
            } else if (td->inputs[i].default_value) {
                args[i] = td->inputs[i].default_value;
            } else {
                // MissingDependency (dst/compute.rs:221-225)
                fill_error(res, AFK_ERR_ARGUMENT, fmt("No output attached to input #%zu of node #%lld in '%s', cannot compute", i, (long long)t, td->name));
                failed = true;
            }
        }
        if (!failed) {
            afk_transform_call(td, args.data(), td->n_inputs, res.data(), res.size());
            g->calls += 1;
            for (afk_result &r : res)
                if (r.status != AFK_OK) {
                    // RuntimeError / ArgumentError display (dst/compute.rs:107-116)
                    char *disp = afk_result_display(&r);
                    std::string m = fmt("%s\n    in node #%lld %s", disp ? disp : "", (long long)t, td->name);
                    free(disp);
                    free(r.message);
                    r.message = strdup(m.c_str());
                }
        }
    }
    clear_cache(n);
    n->cache.values = res;
    n->cache.time = upd;
    n->cache.valid = true;
    *out = &n->cache.values;
    return AFK_OK;
}

extern "C" {

afk_graph *afk_graph_new(void) {
    registry_init();
    return new afk_graph();
}

void afk_graph_free(afk_graph *g) {
    if (!g) return;
    for (Node *n : g->nodes) {
        clear_cache(n);
        if (n->constant) afk_value_release(n->constant);
        delete n;
    }
    delete g;
}

int64_t afk_graph_add_transform(afk_graph *g, const char *name) {
    if (!g) return -1;
    const afk_transform_desc *t = afk_get_transform(name);
    if (!t) {
        fail(make_error(AFK_ERR_UNEXPECTED_INPUT, fmt("Transform '%s' not found", name ? name : "")));
        return -1;
    }
    std::lock_guard<std::mutex> lk(g->mu);
    Node *n = new Node();
    n->t = t;
    n->inputs.resize(t->n_inputs);
    n->updated_on = now();
    g->nodes.push_back(n);
    return (int64_t)g->nodes.size() - 1;
}

int64_t afk_graph_add_constant(afk_graph *g, const afk_value *v) {
    if (!g || !v) return -1;
    std::lock_guard<std::mutex> lk(g->mu);
    Node *n = new Node();
    n->is_constant = true;
    n->constant = const_cast<afk_value *>(v);
    afk_value_retain(n->constant);
    n->updated_on = now();
    g->nodes.push_back(n);
    return (int64_t)g->nodes.size() - 1;
}

int afk_graph_set_constant(afk_graph *g, int64_t t_idx, const afk_value *v) {
    if (!g || !v) return AFK_ERR_UNEXPECTED_INPUT;
    std::lock_guard<std::mutex> lk(g->mu);
    Node *n = node_at(g, t_idx);
    if (!n || !n->is_constant) return fail(make_error(AFK_ERR_UNEXPECTED_INPUT, "not a constant node"));
    afk_value_retain(const_cast<afk_value *>(v));
    afk_value_release(n->constant);
    n->constant = const_cast<afk_value *>(v);
    n->updated_on = now(); // Transform::set_constant -> updated_now()
    return AFK_OK;
}

afk_status afk_graph_connect(afk_graph *g, int64_t from_t, size_t out_i, int64_t to_t, size_t in_i) {
    if (!g) return AFK_ERR_UNEXPECTED_INPUT;
    std::lock_guard<std::mutex> lk(g->mu);
    Node *src = node_at(g, from_t), *dst = node_at(g, to_t);
    // DSTError variants of dst/build.rs:45-88
    if (!src || out_i >= n_outputs(src)) return fail(make_error(AFK_ERR_ARGUMENT, fmt("Output #%zu of node #%lld does not exist in this graph!", out_i, (long long)from_t)));
    if (!dst || dst->is_constant || in_i >= dst->t->n_inputs) return fail(make_error(AFK_ERR_ARGUMENT, fmt("Input #%zu of node #%lld does not exist in this graph!", in_i, (long long)to_t)));
    if (dst->inputs[in_i].from_t == from_t && dst->inputs[in_i].out_i == out_i) return fail(make_error(AFK_ERR_ARGUMENT, "There already is an edge connecting this output to this input!"));
    if (depends_on(g, from_t, to_t)) return fail(make_error(AFK_ERR_ARGUMENT, "Connecting this output to this input would create a cycle!"));
    if (!afk_type_convertible(output_type(src, out_i), dst->t->inputs[in_i].type_id))
        return fail(make_error(AFK_ERR_ARGUMENT, fmt("Cannot connect Output #%zu of node #%lld to Input #%zu of node #%lld. Output does not provide the required input type.",
                                                     out_i, (long long)from_t, in_i, (long long)to_t)));
    dst->inputs[in_i].from_t = from_t; // an input has at most one producer: the previous edge is replaced
    dst->inputs[in_i].out_i = out_i;
    dst->updated_on = now();
    return AFK_OK;
}

afk_status afk_graph_disconnect(afk_graph *g, int64_t to_t, size_t in_i) {
    if (!g) return AFK_ERR_UNEXPECTED_INPUT;
    std::lock_guard<std::mutex> lk(g->mu);
    Node *dst = node_at(g, to_t);
    if (!dst || dst->is_constant || in_i >= dst->t->n_inputs) return fail(make_error(AFK_ERR_ARGUMENT, "input does not exist in this graph"));
    dst->inputs[in_i].from_t = -1;
    dst->updated_on = now();
    return AFK_OK;
}

afk_status afk_graph_compute(afk_graph *g, int64_t t_idx, size_t out_i, afk_result *out) {
    if (!g || !out) return AFK_ERR_UNEXPECTED_INPUT;
    *out = afk_result{AFK_OK, nullptr, nullptr};
    std::lock_guard<std::mutex> lk(g->mu);
    Node *n = node_at(g, t_idx);
    if (!n) {
        Error e = make_error(AFK_ERR_ARGUMENT, fmt("Node #%lld not found", (long long)t_idx));
        out->status = e.status;
        out->message = strdup(e.message.c_str());
        return fail(e);
    }
    if (out_i >= n_outputs(n)) {
        Error e = make_error(AFK_ERR_ARGUMENT, fmt("Output #%zu of node #%lld is not found", out_i, (long long)t_idx));
        out->status = e.status;
        out->message = strdup(e.message.c_str());
        return fail(e);
    }
    std::vector<afk_result> *res = nullptr;
    afk_status st = compute_node(g, t_idx, &res);
    if (st != AFK_OK) {
        out->status = st;
        out->message = strdup(afk_last_error());
        return st;
    }
    copy_result((*res)[out_i], out);
    if (out->status != AFK_OK && out->message) set_last_error(out->message);
    return out->status;
}

uint64_t afk_graph_calls(const afk_graph *g) { return g ? g->calls : 0; }
void afk_graph_set_fusion(afk_graph *g, int enable) {
    if (g) g->fusion = enable != 0;
}

} // extern "C"
