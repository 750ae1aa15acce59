// isect_shadow.cu — the shadow-query instantiations of k_isect3 (MODE 2: any hit towards light j, reference
// trace.py:86-96).  A separate translation unit = a separate CUDA module with its own constant bank: the lights'
// pivot tables stay resident here while the closest-hit tables stay resident in b200rt.cu's bank (see isect3.cuh).
#include "isect3.cuh"
#include <vector>

namespace b200rt {

cudaError_t shadow_set_attr() { return isect3_set_attr<2>(); }

unsigned long long shadow_check_failures() {            // RT_CHECK failures counted in this translation unit (checked build only)
#ifdef B200RT_CHECK
    unsigned int h[8] = {}; unsigned long long t = 0;
    if (cudaMemcpyFromSymbol(h, g_check_fail, sizeof(h)) == cudaSuccess) for (unsigned int v : h) t += v;
    return t;
#else
    return 0;
#endif
}

cudaError_t shadow_upload_resident(const PivotDev& pv, int Spad, int L, cudaStream_t st) {
    std::vector<int> tabs(L);
    for (int j = 0; j < L; ++j) tabs[j] = j;                 // slot j = light j
    return isect3_upload_resident(pv, Spad, tabs.data(), L, st);
}

cudaError_t shadow_launch_light(const IsectCfg& cfg, const SceneDev& sc, const QueueDev& q, const PivotDev& pv, size_t max_rays,
                                const ShadowPlan& sp, int j, bool resident, cudaStream_t st, unsigned long long* tlbase, int* tlnext) {
    return isect3_launch_shadow_light<2>(cfg, sc, q, pv, max_rays, sp, j, resident, st, tlbase, tlnext);
}

}  // namespace b200rt
