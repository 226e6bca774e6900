// solver.cuh -- placeholder, replaced by the LM solver.
#pragma once
#include "handle.cuh"
namespace mvba {
struct Solver {};
inline void solver_destroy(mvba_problem* h) { delete h->solver; h->solver = nullptr; }
inline int solver_solve(mvba_problem*, const mvba_solve_options&, mvba_summary*, mvba_iteration*, int) { return fail(MVBA_ESTATE, "solver not built"); }
inline int solver_bench_build(mvba_problem*, int, float*) { return fail(MVBA_ESTATE, "solver not built"); }
inline int comm_unique_id(uint8_t*) { return fail(MVBA_ENCCL, "nccl not built"); }
inline int comm_init(mvba_problem*, int, int, const uint8_t*) { return fail(MVBA_ENCCL, "nccl not built"); }
}
