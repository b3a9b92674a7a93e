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
int comm_vmm_export(b2nn_net *net, int32_t *pid, int32_t *fd);
int comm_vmm_import(b2nn_net *net, const int32_t *pids, const int32_t *fds, int world);
int comm_mc_create(b2nn_net *net, int32_t *pid, int32_t *fd);
int comm_mc_import(b2nn_net *net, int32_t pid, int32_t fd);
int comm_mc_add_device(b2nn_net *net);
int comm_mc_bind(b2nn_net *net);
void comm_vmm_destroy(b2nn_net *net);
void comm_destroy(b2nn_net *net);
int comm_allreduce_sum_f32(b2nn_net *net, float *buf, size_t count, cudaStream_t st);
}  // namespace b2nn
