// tests/cpp/orbm_oracle_backend.cpp — TEST INFRASTRUCTURE ONLY. Never part of liborbx.so, never loaded by the package.
//
// Purpose: pin the oracle's search restatements (oracle/orb_oracle.cpp, orc_*) AND the host halves of the drop-in
// templates (include/dropin/ORBmatcher.h: projection gates, flattening, write-back, map mutations) against the
// unmodified reference on a machine WITHOUT a GPU. tests/cpp/scene_run.cpp is linked with this file instead of
// liborbx.so: every orbm_* search entry point the templates call is answered by the oracle function the GPU tests use as
// their checker, so `scene_run_hostcheck(scene) == scene_run_ref(scene)` proves, on the CPU, that template host half +
// orc_* == src/ORBmatcher.cc. The GPU tests then run the very same scenes through the real liborbx.so.
// The product has no such path: liborbx.so fails loudly without a device (tests/test_cpu_abi.py).
#include <cstdint>
#include <cstring>
#include <vector>

#include "orbx.h"

struct OrcProj;  // layout-identical to orbm_proj_params (oracle/orb_oracle.cpp)
extern "C" {
int orc_hamming(const uint8_t* a, const uint8_t* b);
int orc_search_by_projection(const OrcProj* Pp, int nLast, const float* worldPos, const uint8_t* lflags,
                             const int32_t* loct, const float* langle, const uint8_t* ldesc, int nCur, const float* kx,
                             const float* ky, const int32_t* koct, const float* kangle, const float* uRight,
                             const uint8_t* cdesc, const uint8_t* taken, int32_t* curMatch, int32_t* lastMatch,
                             uint8_t* lastKeep);
int orc_search_local_points(const OrcProj* Pp, float nnRatio, int nMP, const float* projX, const float* projY,
                            const float* projXR, const int32_t* level, const float* viewCos, const uint8_t* flags,
                            const uint8_t* mdesc, int nCur, const float* kx, const float* ky, const int32_t* koct,
                            const float* uRight, const uint8_t* cdesc, const uint8_t* taken, int32_t* curMatch,
                            int32_t* mpMatch);
int orc_search_by_bow(int nA, const uint8_t* eligA, const float* angA, const uint8_t* descA, int nNodesA,
                      const uint32_t* nodeIdA, const int32_t* offA, const uint32_t* featA, int nB, const uint8_t* eligB,
                      const float* angB, const uint8_t* descB, int nNodesB, const uint32_t* nodeIdB, const int32_t* offB,
                      const uint32_t* featB, float nnRatio, int th, int strict, int checkOri, int32_t* matchA);
int orc_search_for_triangulation(int nA, const uint8_t* eligA, const float* angA, const uint8_t* descA, int nNodesA,
                                 const uint32_t* nodeIdA, const int32_t* offA, const uint32_t* featA, int nB,
                                 const uint8_t* eligB, const float* angB, const uint8_t* descB, int nNodesB,
                                 const uint32_t* nodeIdB, const int32_t* offB, const uint32_t* featB, const float* ptA,
                                 const float* ptB, const int32_t* octB, const uint8_t* stereoA, const uint8_t* stereoB,
                                 const float* F12, float ex, float ey, const float* scaleFactors, const float* levelSigma2,
                                 int th, int checkOri, int32_t* matchA);
void orc_fuse_search(const OrcProj* Pp, const float* invSigma2, int chi2, int n, const float* pu, const float* pv,
                     const float* pur, const int32_t* level, const uint8_t* valid, const uint8_t* desc, int nK,
                     const float* kx, const float* ky, const int32_t* koct, const float* uRight, const uint8_t* kdesc,
                     int32_t* bestIdxOut, uint16_t* bestDistOut);
int orc_search_kf_projection(const OrcProj* Pp, int n, const float* pu, const float* pv, const int32_t* level,
                             const uint8_t* valid, const uint8_t* desc, int nK, const float* kx, const float* ky,
                             const int32_t* koct, const uint8_t* kdesc, const uint8_t* taken, int thLow, int32_t* kfMatch,
                             int32_t* ptMatch);
int orc_search_for_initialization(const OrcProj* Pp, float nnRatio, int thLow, int windowSize, int n1, const int32_t* oct1,
                                  const float* ang1, const uint8_t* desc1, float* prev, int n2, const float* kx,
                                  const float* ky, const int32_t* koct, const float* kang, const uint8_t* desc2,
                                  int32_t* vnMatches12);
void orc_distinctive_descriptors(const uint8_t* desc, const int32_t* off, int npoints, int32_t* bestIdx, int32_t* bestMedian);
}

namespace {
struct Soa {  // keypoint columns the orc_* functions take
    std::vector<float> x, y, ang;
    std::vector<int32_t> oct;
    explicit Soa(const orbx_kp* k, int n) : x(n), y(n), ang(n), oct(n) {
        for (int i = 0; i < n; i++) { x[i] = k[i].x; y[i] = k[i].y; ang[i] = k[i].angle; oct[i] = k[i].octave; }
    }
};
const OrcProj* P(const orbm_proj_params* p) { return reinterpret_cast<const OrcProj*>(p); }
}  // namespace

struct orbm_handle { int unused; };

extern "C" {

int orbm_create(int, orbm_t** out) { *out = new orbm_handle(); return ORBX_OK; }
int orbm_destroy(orbm_t* m) { delete m; return ORBX_OK; }
int orbm_keyframe_hint(orbm_t*, uint64_t) { return ORBX_OK; }
const char* orbx_last_error_string(void) { return "oracle backend"; }

int orbm_distance(orbm_t*, const uint8_t* a, const uint8_t* b, int n, uint16_t* d) {
    for (int i = 0; i < n; i++) d[i] = (uint16_t)orc_hamming(a + 32 * (size_t)i, b + 32 * (size_t)i);
    return ORBX_OK;
}

int orbm_search_by_projection(orbm_t*, const orbm_proj_params* p, const orbm_last_frame* last, const orbm_cur_frame* cur,
                              int32_t* cur_match, int32_t* last_match, uint8_t* last_keep, int* nmatches) {
    Soa c(cur->keys_un, cur->n);
    *nmatches = orc_search_by_projection(P(p), last->n, last->world_pos, last->flags, last->octave, last->angle, last->desc,
                                         cur->n, c.x.data(), c.y.data(), c.oct.data(), c.ang.data(), cur->u_right, cur->desc,
                                         cur->taken, cur_match, last_match, last_keep);
    return ORBX_OK;
}

int orbm_search_local_points(orbm_t*, const orbm_proj_params* p, float nn_ratio, const orbm_map_points* mps,
                             const orbm_cur_frame* cur, int32_t* cur_match, int32_t* mp_match, int* nmatches) {
    Soa c(cur->keys_un, cur->n);
    *nmatches = orc_search_local_points(P(p), nn_ratio, mps->n, mps->proj_x, mps->proj_y, mps->proj_xr, mps->level,
                                        mps->view_cos, mps->flags, mps->desc, cur->n, c.x.data(), c.y.data(), c.oct.data(),
                                        cur->u_right, cur->desc, cur->taken, cur_match, mp_match);
    return ORBX_OK;
}

int orbm_fuse_search(orbm_t*, const orbm_proj_params* p, const float* inv_level_sigma2, int check_chi2,
                     const orbm_fuse_points* pts, const orbm_cur_frame* kf, int32_t* best_idx, uint16_t* best_dist) {
    Soa c(kf->keys_un, kf->n);
    orc_fuse_search(P(p), inv_level_sigma2, check_chi2, pts->n, pts->proj_x, pts->proj_y, pts->proj_xr, pts->level, pts->valid,
                    pts->desc, kf->n, c.x.data(), c.y.data(), c.oct.data(), kf->u_right, kf->desc, best_idx, best_dist);
    return ORBX_OK;
}

int orbm_search_kf_projection(orbm_t*, const orbm_proj_params* p, const orbm_fuse_points* pts, const orbm_cur_frame* kf,
                              int32_t* kf_match, int32_t* pt_match, int* nmatches) {
    Soa c(kf->keys_un, kf->n);
    *nmatches = orc_search_kf_projection(P(p), pts->n, pts->proj_x, pts->proj_y, pts->level, pts->valid, pts->desc, kf->n,
                                         c.x.data(), c.y.data(), c.oct.data(), kf->desc, kf->taken, p->th_high, kf_match,
                                         pt_match);
    return ORBX_OK;
}

int orbm_search_for_initialization(orbm_t*, const orbm_proj_params* p, float nn_ratio, int th_low, int window_size, int n1,
                                   const orbx_kp* keys1, const uint8_t* desc1, float* prev_matched, const orbm_cur_frame* f2,
                                   int32_t* matches12, int* nmatches) {
    Soa a(keys1, n1), c(f2->keys_un, f2->n);
    *nmatches = orc_search_for_initialization(P(p), nn_ratio, th_low, window_size, n1, a.oct.data(), a.ang.data(), desc1,
                                              prev_matched, f2->n, c.x.data(), c.y.data(), c.oct.data(), c.ang.data(),
                                              f2->desc, matches12);
    return ORBX_OK;
}

int orbm_search_by_bow(orbm_t*, const orbm_bow_side* a, const orbm_bow_side* b, int, int, float nn_ratio, int th, int strict,
                       int check_orientation, int32_t* match_a, int* nmatches) {
    std::vector<uint8_t> ones;
    const uint8_t* ea = a->eligible;
    if (!ea) { ones.assign(a->n, 1); ea = ones.data(); }
    *nmatches = orc_search_by_bow(a->n, ea, a->angle, a->desc, a->n_nodes, a->node_id, a->node_off, a->node_feat, b->n,
                                  b->eligible, b->angle, b->desc, b->n_nodes, b->node_id, b->node_off, b->node_feat, nn_ratio,
                                  th, strict, check_orientation, match_a);
    return ORBX_OK;
}

int orbm_search_by_bow_batch(orbm_t* m, int npairs, const orbm_bow_side* a, const orbm_bow_side* b, const int* n_pos_a,
                             const int* n_pos_b, float nn_ratio, int th, int strict, int check_orientation,
                             int32_t* const* match_a, int* nmatches) {
    for (int k = 0; k < npairs; k++)
        orbm_search_by_bow(m, a + k, b + k, n_pos_a[k], n_pos_b[k], nn_ratio, th, strict, check_orientation, match_a[k], nmatches + k);
    return ORBX_OK;
}

int orbm_search_for_triangulation(orbm_t*, const orbm_bow_side* a, const orbm_bow_side* b, int, int,
                                  const orbm_triangulation* t, int th, int check_orientation, int32_t* match_a, int* nmatches) {
    std::vector<uint8_t> oa, ob;
    const uint8_t *ea = a->eligible, *eb = b->eligible;
    if (!ea) { oa.assign(a->n, 1); ea = oa.data(); }
    if (!eb) { ob.assign(b->n, 1); eb = ob.data(); }
    *nmatches = orc_search_for_triangulation(a->n, ea, a->angle, a->desc, a->n_nodes, a->node_id, a->node_off, a->node_feat,
                                             b->n, eb, b->angle, b->desc, b->n_nodes, b->node_id, b->node_off, b->node_feat,
                                             t->pt_a, t->pt_b, t->octave_b, t->stereo_a, t->stereo_b, t->F12, t->ex, t->ey,
                                             t->scale_factors, t->level_sigma2, th, check_orientation, match_a);
    return ORBX_OK;
}

int orbm_distinctive_descriptors(orbm_t*, const uint8_t* desc, const int32_t* off, int npoints, int32_t* best_idx,
                                 int32_t* best_median) {
    std::vector<int32_t> med(npoints > 0 ? npoints : 1);
    orc_distinctive_descriptors(desc, off, npoints, best_idx, best_median ? best_median : med.data());
    return ORBX_OK;
}

}  // extern "C"
