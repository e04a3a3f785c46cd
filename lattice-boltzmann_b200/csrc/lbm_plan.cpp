// lbm_plan.cpp - see lbm_plan.hpp
#include "lbm_plan.hpp"

#include <algorithm>
#include <cstdlib>
#include <set>
#include <sstream>
#include <stdexcept>

namespace lbm {

SlabShape allocate_rows(int ny, int size, int rank)
{
    // d2q9-bgk.c:484-496: minimum_rows = ny / size, the first ny % size ranks get one more.
    SlabShape s;
    const int minimum_rows = ny / size, remainder = ny % size;
    s.rows = minimum_rows + (rank < remainder ? 1 : 0);
    s.row_start = rank * minimum_rows + std::min(rank, remainder);
    return s;
}

int Layout::strip_of(int ch) const
{
    int s = (int)((int64_t)ch * strips / tpr);
    while (s + 1 < strips && c0[s + 1] <= ch) ++s;
    while (s > 0 && c0[s] > ch) --s;
    return s;
}

int64_t Layout::pos(int lr, int ch) const
{
    const int s = strip_of(ch);
    const int w = c0[s + 1] - c0[s];
    return (int64_t)rows * c0[s] + (int64_t)lr * w + (ch - c0[s]);
}

void Layout::task_at(int64_t p, int* lr, int* ch) const
{
    int s = (int)(p / ((int64_t)rows * tpr / strips + 1));
    if (s >= strips) s = strips - 1;
    while (s + 1 < strips && (int64_t)rows * c0[s + 1] <= p) ++s;
    while (s > 0 && (int64_t)rows * c0[s] > p) --s;
    const int w = c0[s + 1] - c0[s];
    const int64_t q = p - (int64_t)rows * c0[s];
    *lr = (int)(q / w);
    *ch = c0[s] + (int)(q % w);
}

int Layout::owner(int lr, int ch) const
{
    const int64_t q = pos(lr, ch);
    return (int)(std::upper_bound(cut.begin(), cut.end(), q) - cut.begin()) - 1;
}

int default_edge_weight_16()
{
    if (const char* e = getenv("LBM_EDGE_WEIGHT_16")) {
        const int v = atoi(e);
        if (v >= 16 && v <= 64) return v;
    }
    // an edge task costs about 1.75 ordinary ones.  Round-1 kernel, 2 GPUs, 16384x128 per GPU: 16 -> 73.4, 20 -> 75.9, 24 -> 79.1,
    // 28 -> 78.5, 32 -> 75.5 GLUPS/GPU.  Round-2 kernel (records pushed except on the last step, 16-warp CTAs on thin slabs), same
    // shape, two passes: 24 -> 79.2, 26 -> 80.0, 28 -> 80.6, 30 -> 79.5, 32 -> 76.3; with 2 x 8-warp CTAs 24 -> 79.3, 28 -> 80.3;
    // 16384x1024 on 8 GPUs: 24 -> 608.8, 28 -> 615.8 (profiles/r02_cta_shape_sweep.md)
    return 28;
}

Layout make_layout(int tpr, int rows, int ctas, bool edge_rows, int edge_weight_16)
{
    // A CTA's patch is ~n/w rows of a strip w chunks wide; its interior is what is not on the rim (a
    // full-width strip has no left/right rim: x is periodic inside the CTA's own rows).  Measured on
    // B200 (profiles/r01_sweep.md): wide strips stream best from HBM, but a patch with no interior
    // (e.g. 16384x128 per GPU at 8 GPUs: half a row per CTA) loses 13 % to flag latency.  So: the FEWEST
    // strips whose expected interior share reaches 30 %, else the count with the largest share.
    Layout best;
    double best_score = -1.0;
    const double n = (double)tpr * rows / ctas;     // tasks per CTA
    for (int strips = 1; strips <= tpr; ++strips) {
        const double w = (double)tpr / strips;
        if (strips > 1 && w < 3.0) break;
        const double h = n / w;
        const double score = std::max(0.0, h - 2.0) * (strips == 1 ? w : std::max(0.0, w - 2.0)) / n;
        if (score > best_score) {
            best_score = score;
            best.strips = strips;
        }
        if (score >= 0.30) break;
    }
    if (const char* e = getenv("LBM_STRIPS")) {       // tuning override
        const int v = atoi(e);
        if (v >= 1 && v <= tpr) best.strips = v;
    }
    best.tpr = tpr; best.rows = rows; best.ctas = ctas;
    best.c0.resize(best.strips + 1);
    for (int s = 0; s <= best.strips; ++s) best.c0[s] = (int)((int64_t)tpr * s / best.strips);

    // cut the sequence into `ctas` ranges of equal weight (every range keeps at least one task)
    const int64_t T = best.tasks();
    const int we = (edge_rows && rows > 2) ? edge_weight_16 : 16;
    std::vector<int64_t> cum(T + 1, 0);
    {
        int64_t q = 0;                            // strip by strip, row-major inside a strip
        for (int s = 0; s < best.strips; ++s) {
            const int w = best.c0[s + 1] - best.c0[s];
            for (int lr = 0; lr < rows; ++lr) {
                const int wt = (edge_rows && (lr == 0 || lr == rows - 1)) ? we : 16;
                for (int i = 0; i < w; ++i, ++q) cum[q + 1] = cum[q] + wt;
            }
        }
    }
    best.cut.assign(ctas + 1, 0);
    for (int b = 1; b < ctas; ++b) {
        const int64_t want = cum[T] * b / ctas;
        int64_t q = std::lower_bound(cum.begin(), cum.end(), want) - cum.begin();
        q = std::max<int64_t>(q, best.cut[b - 1] + 1);
        q = std::min<int64_t>(q, T - (ctas - b));
        best.cut[b] = q;
    }
    best.cut[ctas] = T;
    return best;
}

SlabPlan make_plan(int nx, int ny, int rank, int size, int ctas)
{
    if (nx < 1 || ny < 1 || size < 1 || size > ny || rank < 0 || rank >= size || ctas < 1)
        throw std::runtime_error("make_plan: bad arguments");
    SlabPlan p;
    const SlabShape me = allocate_rows(ny, size, rank);
    p.nx = nx; p.rows = me.rows; p.row_start = me.row_start; p.rank = rank; p.size = size;
    p.tpr = (nx + kTaskCells - 1) / kTaskCells;
    p.num_tasks = p.tpr * p.rows;
    if (ctas > p.num_tasks) ctas = p.num_tasks;
    p.num_ctas = ctas;
    const bool single = (size == 1);

    // global row ny-2 receives the body force (d2q9-bgk.c:241-243, 498)
    const int ar = ny - 2;
    p.accel_local_row = (ar >= me.row_start && ar < me.row_start + me.rows) ? ar - me.row_start : -1;

    const int G = ctas;
    const int64_t T = p.num_tasks;
    p.edge_weight_16 = default_edge_weight_16();
    const Layout L = make_layout(p.tpr, p.rows, G, !single, p.edge_weight_16);
    p.strips = L.strips;

    p.task_begin.assign(G + 1, 0);
    p.n_interior.assign(G, 0);
    p.n_edge.assign(G, 0);
    p.dep_begin.assign(G + 1, 0);
    p.task_order.reserve(T);

    // owner of every task (flat id) and the tasks in sequence order, in one pass over the strips
    std::vector<int> own(T), seq(T);
    {
        int64_t q = 0;
        int b = 0;
        for (int s = 0; s < L.strips; ++s)
            for (int lr = 0; lr < p.rows; ++lr)
                for (int ch = L.c0[s]; ch < L.c0[s + 1]; ++ch, ++q) {
                    while (q >= L.cut[b + 1]) ++b;
                    own[lr * p.tpr + ch] = b;
                    seq[q] = lr * p.tpr + ch;
                }
    }

    std::vector<int> interior, boundary, edge, dep, stamp(G, -1);
    for (int b = 0; b < G; ++b) {
        interior.clear(); boundary.clear(); edge.clear(); dep.clear();
        for (int64_t q = L.cut[b]; q < L.cut[b + 1]; ++q) {
            const int t = seq[q], lr = t / p.tpr, ch = t % p.tpr;
            bool inside = true;
            for (int dy = -1; dy <= 1; ++dy) {
                int r2 = lr + dy;
                if (r2 < 0 || r2 >= p.rows) {
                    if (!single) { inside = false; continue; }   // remote row: tagged halo slots, no flag
                    r2 = (r2 + p.rows) % p.rows;
                }
                for (int dx = -1; dx <= 1; ++dx) {
                    const int c2 = (ch + dx + p.tpr) % p.tpr;
                    const int o = own[r2 * p.tpr + c2];
                    if (o != b) {
                        inside = false;
                        if (stamp[o] != b) { stamp[o] = b; dep.push_back(o); }
                    }
                }
            }
            const bool is_edge = !single && (lr == 0 || lr == p.rows - 1);
            (inside ? interior : is_edge ? edge : boundary).push_back(t);
        }
        std::sort(dep.begin(), dep.end());
        p.n_interior[b] = (int)interior.size();
        p.n_edge[b] = (int)edge.size();
        p.task_order.insert(p.task_order.end(), interior.begin(), interior.end());
        p.task_order.insert(p.task_order.end(), edge.begin(), edge.end());
        p.task_order.insert(p.task_order.end(), boundary.begin(), boundary.end());
        p.task_begin[b + 1] = (int)p.task_order.size();
        p.deps.insert(p.deps.end(), dep.begin(), dep.end());
        p.dep_begin[b + 1] = (int)p.deps.size();
    }
    return p;
}

std::string validate_plan(const SlabPlan& p)
{
    const int G = p.num_ctas;
    const int64_t T = p.num_tasks;
    const Layout L = make_layout(p.tpr, p.rows, G, p.size > 1, p.edge_weight_16);
    std::vector<int> seen(T, 0);
    if ((int)p.task_order.size() != T) return "task_order size";
    for (int b = 0; b < G; ++b) {
        const int64_t lo = L.cut[b], hi = L.cut[b + 1];
        if (hi <= lo) return "empty CTA range";
        if (p.task_begin[b + 1] - p.task_begin[b] != hi - lo) return "task_begin mismatch";
        std::set<int> dep(p.deps.begin() + p.dep_begin[b], p.deps.begin() + p.dep_begin[b + 1]);
        bool seen_plain = false;
        for (int k = p.task_begin[b]; k < p.task_begin[b + 1]; ++k) {
            const int t = p.task_order[k];
            if (t < 0 || t >= T) return "task id out of range";
            const int lr = t / p.tpr, ch = t % p.tpr;
            if (L.owner(lr, ch) != b) return "task outside owner range";
            if (seen[t]++) return "task listed twice";
            const bool is_int = (k - p.task_begin[b]) < p.n_interior[b];
            const bool is_edge = p.size > 1 && (lr == 0 || lr == p.rows - 1);
            if (is_edge && is_int) return "edge task scheduled as interior";
            if (is_edge && seen_plain) return "edge tasks must lead the boundary tasks";
            const int rel = k - p.task_begin[b] - p.n_interior[b];
            if (is_edge != (rel >= 0 && rel < p.n_edge[b])) return "n_edge does not delimit the edge tasks";
            seen_plain = seen_plain || (!is_int && !is_edge);
            bool all_mine = true;
            for (int dy = -1; dy <= 1; ++dy)
                for (int dx = -1; dx <= 1; ++dx) {
                    const int c2 = (ch + dx + p.tpr) % p.tpr;
                    int r2 = lr + dy;
                    if (r2 < 0 || r2 >= p.rows) {
                        if (p.size == 1) r2 = (r2 + p.rows) % p.rows;
                        else { all_mine = false; continue; }
                    }
                    const int o = L.owner(r2, c2);
                    if (o != b) {
                        all_mine = false;
                        if (is_int) return "interior task with foreign neighbour";
                        if (!dep.count(o)) return "missing dependency";
                    }
                }
            if (all_mine && !is_int) return "self-contained task scheduled as boundary";
        }
        for (int d : dep) if (d < 0 || d >= G || d == b) return "bad dependency index";
    }
    for (int64_t t = 0; t < T; ++t) if (seen[t] != 1) return "task missing";
    return "";
}

// Fraction of tasks that are interior (diagnostic for tests / describe()).
double interior_fraction(const SlabPlan& p)
{
    long long n = 0;
    for (int v : p.n_interior) n += v;
    return p.num_tasks ? (double)n / p.num_tasks : 0.0;
}

}  // namespace lbm
