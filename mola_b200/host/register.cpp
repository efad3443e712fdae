// register.cpp — class registration at library load, like mola_metric_maps/src/register.cpp:41-50,
// plus the MAP_DEFINITION_REGISTER names (HashedVoxelPointCloud.cpp:56-58 and siblings) so that a
// YAML/INI `class: mola::HashedVoxelPointCloud` instantiates the GPU map.
#include <mrpt/core/initializer.h>
#include <mrpt/rtti/CObject.h>

#include "GpuMetricMaps.h"

#define MOLA_B200_STR2(x) #x
#define MOLA_B200_STR(x) MOLA_B200_STR2(x)

MAP_DEFINITION_REGISTER(MOLA_B200_STR(MOLA_B200_NAMESPACE) "::HashedVoxelPointCloud,HashedVoxelPointCloud", MOLA_B200_NAMESPACE::HashedVoxelPointCloud)
MAP_DEFINITION_REGISTER(MOLA_B200_STR(MOLA_B200_NAMESPACE) "::SparseVoxelPointCloud,SparseVoxelPointCloud", MOLA_B200_NAMESPACE::SparseVoxelPointCloud)
MAP_DEFINITION_REGISTER(MOLA_B200_STR(MOLA_B200_NAMESPACE) "::SparseTreesPointCloud,SparseTreesPointCloud", MOLA_B200_NAMESPACE::SparseTreesPointCloud)

namespace MOLA_B200_NAMESPACE
{
void registerGpuMetricMaps()
{
    using mrpt::rtti::registerClass;
    registerClass(CLASS_ID(HashedVoxelPointCloud));
    registerClass(CLASS_ID(SparseVoxelPointCloud));
    registerClass(CLASS_ID(SparseTreesPointCloud));
}
}  // namespace MOLA_B200_NAMESPACE

MRPT_INITIALIZER(do_register_mola_b200_metric_maps) { MOLA_B200_NAMESPACE::registerGpuMetricMaps(); }
