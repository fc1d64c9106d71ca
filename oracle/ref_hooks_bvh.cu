// ref_hooks_bvh.cu — compiles the reference's src/bvh.cu as it lies and exposes its host-callable Morton code
// (getMortonCode / floatTo21Int / expandBits, src/bvh.cu:14-62).  Test infrastructure only.
#include "bvh.cu"

extern "C" long long ref_morton(float x, float y, float z) { return (long long)getMortonCode(Vec3(x, y, z)); }
