// oracle/ref_fbow_bridge.cpp — TEST INFRASTRUCTURE ONLY.
// C entry points over the UNMODIFIED reference bag-of-words library fbow::Vocabulary / fbow::fBow / fbow::fBow2
// (/root/reference/Thirdparty/fbow/src/fbow.h:31-450, fbow.cpp), which oracle/Makefile compiles from where it lies
// against oracle/shim. Output: oracle/_ref/libref_fbow.so (git-ignored). Used to validate the restatement of the
// vocabulary-tree descent in orb_oracle.cpp (orc_fbow_*) and the .fbow blob writer in orbslam2_b200/synth.py.
#include <cstdint>
#include <cstring>
#include <sstream>
#include <string>

#include "fbow.h"

extern "C" {

// blob = the bytes of a .fbow file (Vocabulary::toStream format, fbow.cpp:164-183)
void* ref_fbow_load(const char* blob, size_t size) {
    try {
        std::istringstream in(std::string(blob, size), std::ios::binary);
        fbow::Vocabulary* v = new fbow::Vocabulary();
        v->fromStream(in);
        return v;
    } catch (...) {
        return nullptr;
    }
}
void ref_fbow_free(void* h) { delete (fbow::Vocabulary*)h; }

void ref_fbow_info(void* h, uint32_t* k, uint64_t* nblocks, uint32_t* desc_size, uint64_t* total_words) {
    fbow::Vocabulary& v = *(fbow::Vocabulary*)h;
    *k = v.getK();
    *nblocks = v.size();
    *desc_size = v.getDescSize();
    *total_words = v.getTotalWords();
}

static int flatten(const fbow::fBow& r1, uint32_t* wid, float* ww) {
    int n = 0;
    for (const auto& e : r1) {
        wid[n] = e.first;
        ww[n] = (float)e.second;
        n++;
    }
    return n;
}

// Vocabulary::transform(features, level, fBow&, fBow2&) (fbow.cpp:52-93 -> _transform2<L1_32bytes>, fbow.h:400-444).
// Returns -1 when the reference throws. noff has room for n+1 entries, feat for n.
int ref_fbow_transform2(void* h, const uint8_t* desc, int n, int level, uint32_t* wid, float* ww, int* n_words,
                        uint32_t* nid, int* noff, uint32_t* feat, int* n_nodes) {
    fbow::Vocabulary& v = *(fbow::Vocabulary*)h;
    cv::Mat f(n, 32, CV_8UC1, (void*)desc, 32);
    fbow::fBow r1;
    fbow::fBow2 r2;
    try {
        v.transform(f, level, r1, r2);
    } catch (...) {
        return -1;
    }
    *n_words = flatten(r1, wid, ww);
    int nn = 0, at = 0;
    for (const auto& e : r2) {
        nid[nn] = e.first;
        noff[nn] = at;
        for (uint32_t i : e.second) feat[at++] = i;
        nn++;
    }
    noff[nn] = at;
    *n_nodes = nn;
    return 0;
}

// Vocabulary::transform(features) (fbow.cpp:95-153): leaf weights summed per word, then L2-normalised.
int ref_fbow_transform(void* h, const uint8_t* desc, int n, uint32_t* wid, float* ww, int* n_words) {
    fbow::Vocabulary& v = *(fbow::Vocabulary*)h;
    cv::Mat f(n, 32, CV_8UC1, (void*)desc, 32);
    try {
        fbow::fBow r = v.transform(f);
        *n_words = flatten(r, wid, ww);
    } catch (...) {
        return -1;
    }
    return 0;
}

// fBow::score (fbow.cpp:208-254)
double ref_fbow_score(const uint32_t* id1, const float* w1, int n1, const uint32_t* id2, const float* w2, int n2) {
    fbow::fBow a, b;
    for (int i = 0; i < n1; i++) { float t = w1[i]; a[id1[i]] = t; }
    for (int i = 0; i < n2; i++) { float t = w2[i]; b[id2[i]] = t; }
    return fbow::fBow::score(a, b);
}

}  // extern "C"
