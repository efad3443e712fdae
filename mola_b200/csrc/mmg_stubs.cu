// Temporary stubs (replaced by mmg_sv.cu / mmg_trees.cu / mmg_nccl.cu as they land).
#include "mmg_host.h"
using namespace mmg;
#define UNSUP(name) { set_error(name ": not implemented yet"); return MMG_ERR_UNSUPPORTED; }
int trees_create(mmg_map*) { return MMG_OK; }
int trees_destroy(mmg_map*) { return MMG_OK; }
int trees_clear(mmg_map*) UNSUP("trees")
int trees_set_grid(mmg_map*) UNSUP("trees")
int trees_clone(const mmg_map*, mmg_map*) UNSUP("trees")
int trees_insert(mmg_map*, const float*, const float*, const float*, size_t, const double*, int, int, cudaStream_t) UNSUP("trees")
int trees_nn_single(const mmg_map*, const float*, const float*, const float*, size_t, size_t, const double*, uint8_t*, float*, float*, uint64_t*, int, cudaStream_t) UNSUP("trees")
int trees_nn_multi(const mmg_map*, const float*, const float*, const float*, size_t, const double*, size_t, uint32_t*, float*, float*, uint64_t*, int, cudaStream_t) UNSUP("trees")
int trees_radius(mmg_map*, const float*, const float*, const float*, size_t, const double*, float, size_t, uint64_t*, uint64_t*, int, cudaStream_t) UNSUP("trees")
int trees_is_empty(const mmg_map*, int* o) { *o = 1; return MMG_OK; }
int trees_num_blocks(const mmg_map*, uint64_t* o) { *o = 0; return MMG_OK; }
int trees_num_points(const mmg_map*, uint64_t* o) { *o = 0; return MMG_OK; }
int trees_row_capacity(const mmg_map*, uint32_t* o) { *o = 0; return MMG_OK; }
int trees_export(const mmg_map*, int32_t*, uint32_t*, float*, int, cudaStream_t) UNSUP("trees")
int trees_bbox_from_points(mmg_map*) UNSUP("trees")
int trees_table_stats(mmg_map*, mmg_table_stats*) UNSUP("trees")
namespace mmg {
int trees_bcast(mmg_map*, int, int, void*, cudaStream_t, u64, u64, int (*)(void*, void*, size_t, int, cudaStream_t)) UNSUP("trees")
}
