// lbm_plan.cpp - see lbm_plan.hpp
#include "lbm_plan.hpp"

#include <algorithm>
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

int owner_of(int64_t tasks, int ctas, int64_t t)
{
    // smallest b with begin(b+1) > t ; begin(b) = tasks*b/ctas (floor)
    int64_t b = (t * ctas) / tasks;
    while (b + 1 < ctas && range_begin(tasks, ctas, (int)(b + 1)) <= t) ++b;
    while (b > 0 && range_begin(tasks, ctas, (int)b) > t) --b;
    return (int)b;
}

int choose_ctas(int nx, int rows, int max_resident)
{
    const int tpr = (nx + kTaskCells - 1) / kTaskCells;
    const int64_t tasks = (int64_t)tpr * rows;
    int64_t g = std::min<int64_t>(max_resident, tasks);
    return (int)std::max<int64_t>(g, 1);
}

namespace {

// CTAs (ascending) that own at least one task of local row `lr` of a slab.
std::vector<int> row_owners(int tpr, int rows, int ctas, int lr)
{
    const int64_t tasks = (int64_t)tpr * rows;
    const int first = owner_of(tasks, ctas, (int64_t)lr * tpr);
    const int last = owner_of(tasks, ctas, (int64_t)lr * tpr + tpr - 1);
    std::vector<int> v;
    for (int b = first; b <= last; ++b) v.push_back(b);
    return v;
}

int index_in(const std::vector<int>& v, int x)
{
    auto it = std::lower_bound(v.begin(), v.end(), x);
    if (it == v.end() || *it != x) return -1;
    return (int)(it - v.begin());
}

}  // namespace

SlabPlan make_plan(int nx, int ny, int rank, int size, int ctas, int ctas_up, int ctas_down)
{
    if (nx < 1 || ny < 1 || size < 1 || size > ny || rank < 0 || rank >= size || ctas < 1)
        throw std::runtime_error("make_plan: bad arguments");
    SlabPlan p;
    const SlabShape me = allocate_rows(ny, size, rank);
    const int up = (rank + 1) % size, down = (rank - 1 + size) % size;   // d2q9-bgk.c:501-502
    p.nx = nx; p.rows = me.rows; p.row_start = me.row_start; p.rank = rank; p.size = size;
    p.rows_up = allocate_rows(ny, size, up).rows;
    p.rows_down = allocate_rows(ny, size, down).rows;
    p.tpr = (nx + kTaskCells - 1) / kTaskCells;
    p.num_tasks = p.tpr * p.rows;
    if (ctas > p.num_tasks) ctas = p.num_tasks;
    p.num_ctas = ctas;
    const bool single = (size == 1);
    p.ctas_up = single ? ctas : std::min(ctas_up, p.tpr * p.rows_up);
    p.ctas_down = single ? ctas : std::min(ctas_down, p.tpr * p.rows_down);
    if (!single && (p.ctas_up < 1 || p.ctas_down < 1)) throw std::runtime_error("make_plan: bad neighbour ctas");

    // global row ny-2 receives the body force (d2q9-bgk.c:241-243, 498)
    const int ar = ny - 2;
    p.accel_local_row = (ar >= me.row_start && ar < me.row_start + me.rows) ? ar - me.row_start : -1;

    const int G = ctas;
    const int64_t T = p.num_tasks;
    if (!single) {
        p.up_owners = row_owners(p.tpr, p.rows_up, p.ctas_up, 0);
        p.down_owners = row_owners(p.tpr, p.rows_down, p.ctas_down, p.rows_down - 1);
    }
    p.num_flags = G + (int)p.up_owners.size() + (int)p.down_owners.size();
    const int up_base = G, down_base = G + (int)p.up_owners.size();

    p.task_begin.assign(G + 1, 0);
    p.n_interior.assign(G, 0);
    p.dep_begin.assign(G + 1, 0);
    p.task_order.reserve(T);
    p.pub_up.assign(G, -1);
    p.pub_down.assign(G, -1);

    for (int b = 0; b < G; ++b) {
        const int64_t lo = range_begin(T, G, b), hi = range_begin(T, G, b + 1);
        std::vector<int> interior, boundary;
        std::set<int> dep;
        for (int64_t t = lo; t < hi; ++t) {
            const int lr = (int)(t / p.tpr), ch = (int)(t % p.tpr);
            bool inside = true;
            for (int dy = -1; dy <= 1; ++dy) {
                for (int dx = -1; dx <= 1; ++dx) {
                    const int c2 = (ch + dx + p.tpr) % p.tpr;
                    int r2 = lr + dy;
                    if (r2 < 0 || r2 >= p.rows) {
                        if (single) {
                            r2 = (r2 + p.rows) % p.rows;
                        } else {
                            // remote row: last real row of `down` (dy<0) or first real row of `up`
                            inside = false;
                            if (r2 < 0) {
                                const int o = owner_of((int64_t)p.tpr * p.rows_down, p.ctas_down,
                                                       (int64_t)(p.rows_down - 1) * p.tpr + c2);
                                dep.insert(down_base + index_in(p.down_owners, o));
                            } else {
                                const int o = owner_of((int64_t)p.tpr * p.rows_up, p.ctas_up, c2);
                                dep.insert(up_base + index_in(p.up_owners, o));
                            }
                            continue;
                        }
                    }
                    const int o = owner_of(T, G, (int64_t)r2 * p.tpr + c2);
                    if (o != b) { inside = false; dep.insert(o); }
                }
            }
            (inside ? interior : boundary).push_back((int)t);
        }
        p.n_interior[b] = (int)interior.size();
        p.task_order.insert(p.task_order.end(), interior.begin(), interior.end());
        p.task_order.insert(p.task_order.end(), boundary.begin(), boundary.end());
        p.task_begin[b + 1] = (int)p.task_order.size();
        p.deps.insert(p.deps.end(), dep.begin(), dep.end());
        p.dep_begin[b + 1] = (int)p.deps.size();

        if (!single) {
            // I mirror my progress to the DOWN neighbour if I own part of my first real row:
            // there I appear in its up_owners list (its up neighbour is me).
            const int64_t row0_end = p.tpr;                       // tasks [0, tpr)
            const int64_t rowL_begin = (int64_t)(p.rows - 1) * p.tpr;
            if (lo < row0_end) {
                const std::vector<int> mine = row_owners(p.tpr, p.rows, G, 0);
                p.pub_down[b] = p.ctas_down + index_in(mine, b);
            }
            if (hi > rowL_begin) {
                // in the UP neighbour's flag array: after its G and its up_owners section.
                // Its up_owners section length = owners of the first real row of ITS up neighbour.
                const int upup = (up + 1) % size;
                const int rows_upup = allocate_rows(ny, size, upup).rows;
                // all slabs are assumed to use the same CTA budget as passed for `up`
                const int ctas_upup = std::min((upup == rank) ? G : ctas_up, p.tpr * rows_upup);
                const int n_up_section = (int)row_owners(p.tpr, rows_upup, ctas_upup, 0).size();
                const std::vector<int> mine = row_owners(p.tpr, p.rows, G, p.rows - 1);
                p.pub_up[b] = p.ctas_up + n_up_section + index_in(mine, b);
            }
        }
    }
    return p;
}

std::string validate_plan(const SlabPlan& p)
{
    std::ostringstream err;
    const int G = p.num_ctas;
    const int64_t T = p.num_tasks;
    std::vector<int> seen(T, 0);
    if ((int)p.task_order.size() != T) return "task_order size";
    for (int b = 0; b < G; ++b) {
        const int64_t lo = range_begin(T, G, b), hi = range_begin(T, G, b + 1);
        if (p.task_begin[b + 1] - p.task_begin[b] != hi - lo) return "task_begin mismatch";
        std::set<int> dep(p.deps.begin() + p.dep_begin[b], p.deps.begin() + p.dep_begin[b + 1]);
        for (int k = p.task_begin[b]; k < p.task_begin[b + 1]; ++k) {
            const int t = p.task_order[k];
            if (t < lo || t >= hi) return "task outside owner range";
            if (seen[t]++) return "task listed twice";
            const bool is_int = (k - p.task_begin[b]) < p.n_interior[b];
            const int lr = t / p.tpr, ch = t % p.tpr;
            for (int dy = -1; dy <= 1; ++dy)
                for (int dx = -1; dx <= 1; ++dx) {
                    const int c2 = (ch + dx + p.tpr) % p.tpr;
                    int r2 = lr + dy;
                    if (r2 < 0 || r2 >= p.rows) {
                        if (p.size == 1) r2 = (r2 + p.rows) % p.rows;
                        else { if (is_int) return "interior task touches a remote row"; continue; }
                    }
                    const int o = owner_of(T, G, (int64_t)r2 * p.tpr + c2);
                    if (o != b) {
                        if (is_int) return "interior task with foreign neighbour";
                        if (!dep.count(o)) return "missing dependency";
                    }
                }
        }
        for (int d : dep) if (d < 0 || d >= p.num_flags || d == b) return "bad dependency index";
    }
    for (int64_t t = 0; t < T; ++t) if (seen[t] != 1) return "task missing";
    if (p.num_flags != G + (int)p.up_owners.size() + (int)p.down_owners.size()) return "num_flags";
    return "";
}

}  // namespace lbm
