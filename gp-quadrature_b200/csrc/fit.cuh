// Fused posterior-mean fit and the library communicator -- internal declarations (fit.cu, comm.cpp).
#pragma once
#include "common.cuh"

namespace efgp {

struct Nufft;
struct Toeplitz;

// mirrors efgp_fit_args of include/efgp_b200.h field for field
struct FitArgs {
  int family;           // 0: S = c exp(-a |xi|^2);  1: S = c (b + 4 pi^2 |xi|^2)^(-p)
  double c, a, b, p;
  double h, hd;         // grid spacing and h^d (as the host computed it in the model's dtype)
  int mtot;
  double sigmasq, nufft_eps, gen_eps, cg_tol;
  int max_iter, precondition, force_pow2, sharded, warm_start;
};

void bbox_async(const void* x, int64_t N, int dim, int dtype, int sharded, double* ext_pinned, cudaStream_t st);
void fit_run(Nufft& nu, Toeplitz& t, const void* y, const FitArgs& a, void* ws, void* vr, void* minv, void* beta,
             int* counters_pinned, int* iters_completed, cudaStream_t st);

void comm_unique_id(void* id128);
void comm_init(int rank, int world, const void* id128);
void comm_destroy();
int comm_world();
int comm_rank();
void comm_allreduce(void* buf, int64_t count, int dtype, int op, cudaStream_t st);

}  // namespace efgp
