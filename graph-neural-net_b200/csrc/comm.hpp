// comm.hpp -- NCCL plumbing for data-parallel training (one process per GPU).  NCCL is bound at
// run time with dlopen so that libb2nn.so loads on machines without it and shares the copy the
// launcher (torch.distributed) already mapped.
#pragma once
#include "common.hpp"

namespace b2nn {
int comm_unique_id(void *id128);
int comm_init(b2nn_net *net, const void *id128, int rank, int world);
int comm_ipc_export(b2nn_net *net, void *handle64);
int comm_ipc_import(b2nn_net *net, const void *handles, int world);
void comm_destroy(b2nn_net *net);
int comm_allreduce_sum_f32(b2nn_net *net, float *buf, size_t count, cudaStream_t st);
}  // namespace b2nn
