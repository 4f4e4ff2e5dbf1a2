// development aid: test_node() alone, to count its SASS by pipe (tools/scratch is not part of the product)
#include "../../visionaray_b200/csrc/bvh8_traverse.cuh"
using namespace vsb;
__global__ void node_only(const Node80* nodes, const float* rays, uint32_t* out, float limit)
{
    RayCtx r;
    const float* q = rays + 8 * threadIdx.x;
    ray_ctx_init(r, q[0], q[1], q[2], q[4], q[5], q[6]);
    NodeRegs n;
    load_record(reinterpret_cast<const char*>(nodes + threadIdx.x), 5, n);
    uint32_t ph;
    const uint32_t nh = test_node(n, r, limit, ph);
    out[2 * threadIdx.x] = nh; out[2 * threadIdx.x + 1] = ph;
}
