// lbm_plan.hpp - host-side decomposition of the lattice into slabs (one per GPU),
// warp-tasks and per-CTA schedules for the persistent neighbour-synchronised engine.
// Pure C++ (no CUDA): unit-tested on the CPU through the C ABI hooks in lbm_api.cu.
//
// Vocabulary
//   slab      contiguous block of rows owned by one GPU (the reference's per-rank rows,
//             allocate_rows, d2q9-bgk.c:483-503), stored with one halo row either side.
//   task      LBM_TASK_CELLS consecutive cells of one row = what one warp updates at a
//             time (each lane a float4 of 4 cells).  Flat id = local_row * tpr + chunk.
//   CTA range each persistent CTA owns a contiguous range of the strip-major task sequence.
//   interior  a task whose whole 3x3 task neighbourhood (periodic in x; periodic in y only
//             for a single slab) is owned by the same CTA: it can be updated for step s+1
//             as soon as the CTA itself has finished step s.
//   boundary  every other task: needs the neighbouring CTAs to have finished step s, both to read
//             their results and to be sure they no longer read the cells this task overwrites
//             (A/B buffers).
//   edge      a boundary task on the first or last row of a slab that has ring neighbours: the
//             row beyond the edge comes from the neighbouring GPU through tagged halo slots
//             (lbm_device.cuh), which carry their own readiness - no cross-GPU flag.  Edge tasks
//             lead a CTA's boundary tasks: their pushes then have the rest of the step to cross
//             NVLink before the neighbour needs them, and are acknowledged before the CTA releases
//             its progress flag (a release waits for outstanding remote writes).
//   flags     per slab: progress of its CTAs.
#pragma once
#include <cstdint>
#include <string>
#include <vector>

namespace lbm {

constexpr int kTaskCells = 128;   // 32 lanes x float4

struct SlabShape {
    int rows = 0;        // real rows of this slab
    int row_start = 0;   // first global row
};

// The reference's row decomposition (allocate_rows, d2q9-bgk.c:484-496):
// ny/size rows each, the first ny%size ranks one more.
SlabShape allocate_rows(int ny, int size, int rank);

// Ownership layout of one slab: the rows x tpr grid of tasks is cut into vertical STRIPS of about
// `strip_w` chunks; tasks are ordered strip by strip, row-major inside a strip, and that sequence is cut
// into `ctas` equal ranges.  Every CTA therefore owns a compact patch (a few rows of one strip) with a
// perfectly balanced task count, and most of its tasks are interior.  strip_w == tpr is plain row-major.
struct Layout {
    int tpr = 0, rows = 0, ctas = 0, strips = 1;
    std::vector<int> c0;      // first chunk of strip s (size strips+1)
    std::vector<int64_t> cut; // CTA b owns sequence positions [cut[b], cut[b+1])  (size ctas+1)
    int64_t tasks() const { return (int64_t)tpr * rows; }
    int strip_of(int ch) const;
    int64_t pos(int lr, int ch) const;        // position in the strip-major sequence
    void task_at(int64_t pos, int* lr, int* ch) const;
    int owner(int lr, int ch) const;
};
// Picks the strip count that maximises the expected number of interior tasks per CTA and cuts the
// sequence into `ctas` ranges of equal WEIGHT: a task counts 1, except that on a slab with ring
// neighbours (edge_rows) the tasks of the first and last row count edge_weight_16 / 16 - they pull and
// push tagged halo records on top of the ordinary update, and the slowest CTA paces its neighbours.
Layout make_layout(int tpr, int rows, int ctas, bool edge_rows = false, int edge_weight_16 = 16);
int default_edge_weight_16();   // measured on B200 (profiles/r01_sync_probe.md); env LBM_EDGE_WEIGHT_16 overrides

struct SlabPlan {
    int nx = 0, rows = 0, row_start = 0;
    int strips = 1, edge_weight_16 = 16;
    int rank = 0, size = 1;
    int tpr = 0;          // tasks per row
    int num_tasks = 0;
    int num_ctas = 0;     // G
    // schedule (CSR style)
    std::vector<int> task_begin;   // G+1 offsets into task_order
    std::vector<int> n_interior;   // G: the first n_interior[b] tasks of CTA b are interior
    std::vector<int> n_edge;       // G: the next n_edge[b] tasks are slab-edge tasks (0 on a single lattice)
    std::vector<int> task_order;   // num_tasks flat task ids; per CTA: interior, edge, boundary
    std::vector<int> dep_begin;    // G+1 offsets into deps
    std::vector<int> deps;         // CTAs (flag indices) a CTA's boundary tasks wait for
    int accel_local_row = -1;      // local real row that is global row ny-2, else -1
};

// Builds the plan of slab `rank` of `size` for an nx x ny lattice with `ctas` persistent CTAs.
// Throws std::runtime_error on impossible arguments.
SlabPlan make_plan(int nx, int ny, int rank, int size, int ctas);

// Self-check used by the tests: every task appears exactly once, interior tasks really have
// all neighbours inside the CTA, every neighbour owner of a boundary task is in deps, edge tasks
// lead the boundary tasks.  Returns "" or a description.
std::string validate_plan(const SlabPlan& p);
double interior_fraction(const SlabPlan& p);

}  // namespace lbm
