// include/dropin/fbow.h — header-only C++ drop-in for the parts of fbow::Vocabulary / fBow / fBow2 that ORB-SLAM2 calls,
// over the liborbx C ABI (orbv_* in orbx.h).
//
// Put this directory ahead of the reference's Thirdparty/fbow/include on the include path (the reference includes
// "Thirdparty/fbow/include/fbow/fbow.h", include/Frame.h:32) and link liborbx.so. Replaces
// /root/reference/Thirdparty/fbow/src/fbow.h:14-125 and fbow.cpp:52-254 for:
//   Vocabulary::readFromFile / fromStream            src/System.cc:71-80 (vocabulary load)
//   Vocabulary::transform(features, level, fBow&, fBow2&)   src/Frame.cc:395-400, src/KeyFrame.cc:62
//   Vocabulary::transform(features)                  (normalised bag of words)
//   fBow::score                                      src/KeyFrameDatabase.cc:130,247, src/LoopClosing.cc:170
// The tree descent, the per-word weight sums and the scores run in CUDA kernels; this file only marshals std::map
// to and from flat arrays. There is no CPU path: errors from the library become std::runtime_error, as fbow throws.
#ifndef ORBX_DROPIN_FBOW_H
#define ORBX_DROPIN_FBOW_H

#include <cstdint>
#include <cstring>
#include <fstream>
#include <istream>
#include <iterator>
#include <map>
#include <stdexcept>
#include <string>
#include <vector>

#include <opencv2/core/core.hpp>

#include "../orbx.h"

namespace fbow {

// float initialised to zero (fbow.h:14-26)
struct _float {
    float var = 0;
    inline float operator=(float& f) { var = f; return var; }
    inline operator float&() { return var; }
    inline operator float() const { return var; }
};

struct fBow : std::map<uint32_t, _float> {
    typedef std::map<uint32_t, _float> super;
    // fbow.cpp:208-254, evaluated on the device (one pair = a batch of one)
    static double score(const fBow& v1, const fBow& v2) {
        std::vector<const fBow*> db(1, &v2);
        std::vector<double> s;
        scoreBatch(nullptr, v1, db, s);
        return s[0];
    }
    // score(q, *db[i]) for all i in one launch: the loop of KeyFrameDatabase::DetectLoopCandidates (:117-139)
    static void scoreBatch(orbv_t* voc, const fBow& q, const std::vector<const fBow*>& db, std::vector<double>& out) {
        std::vector<uint32_t> qi, di;
        std::vector<float> qw, dw;
        std::vector<int> off(1, 0);
        flatten(q, qi, qw);
        for (const fBow* v : db) {
            flatten(*v, di, dw);
            off.push_back((int)di.size());
        }
        out.assign(db.size(), 0.0);
        if (orbv_score_batch(voc, qi.data(), qw.data(), (int)qi.size(), di.data(), dw.data(), off.data(), (int)db.size(),
                             out.data()) != ORBX_OK)
            throw std::runtime_error(std::string("fBow::score: ") + orbx_last_error_string());
    }
    static void flatten(const fBow& v, std::vector<uint32_t>& id, std::vector<float>& w) {
        for (const auto& e : v) {
            id.push_back(e.first);
            w.push_back((float)e.second);
        }
    }
};

struct fBow2 : std::map<uint32_t, std::vector<uint32_t>> {
    typedef std::map<uint32_t, std::vector<uint32_t>> super;
};

class Vocabulary {
public:
    explicit Vocabulary(int cudaDevice = 0) : device_(cudaDevice) {}
    ~Vocabulary() { clear(); }
    Vocabulary(const Vocabulary&) = delete;
    Vocabulary& operator=(const Vocabulary&) = delete;

    void readFromFile(const std::string& filepath) {
        std::ifstream file(filepath, std::ios::binary);
        if (!file) throw std::runtime_error("Vocabulary::readFromFile could not open:" + filepath);
        fromStream(file);
    }
    void fromStream(std::istream& str) {
        std::vector<char> blob((std::istreambuf_iterator<char>(str)), std::istreambuf_iterator<char>());
        clear();
        if (orbv_create(blob.data(), blob.size(), device_, &h_) != ORBX_OK) {
            h_ = nullptr;
            throw std::runtime_error(std::string("Vocabulary::fromStream ") + orbx_last_error_string());
        }
    }
    void clear() {
        if (h_) orbv_destroy(h_);
        h_ = nullptr;
    }
    bool isValid() const { return h_ != nullptr; }
    uint32_t getK() const { return info(0); }
    size_t size() const { return info(1); }
    uint32_t getDescSize() const { return info(2); }
    uint32_t getDescType() const { return CV_8UC1; }
    size_t getTotalWords() {
        uint64_t w = 0;
        if (h_) orbv_info(h_, nullptr, nullptr, nullptr, &w);
        return (size_t)w;
    }
    orbv_t* handle() const { return h_; }

    // fbow.cpp:52-93
    void transform(const cv::Mat& features, int level, fBow& result, fBow2& result2) {
        run(features, level, 0, result, &result2);
    }
    // fbow.cpp:95-153
    fBow transform(const cv::Mat& features) {
        fBow r;
        run(features, -1, 1, r, nullptr);
        return r;
    }

private:
    uint32_t info(int which) const {
        uint32_t v[3] = {0, 0, 0};
        if (h_) orbv_info(h_, &v[0], &v[1], &v[2], nullptr);
        return v[which];
    }
    void run(const cv::Mat& features, int level, int normalize, fBow& r1, fBow2* r2) {
        if (features.rows == 0) throw std::runtime_error("Vocabulary::transform No input data");
        if (features.type() != CV_8UC1)
            throw std::runtime_error("Vocabulary::transform features are of different type than vocabulary");
        if (!h_ || features.cols * features.elemSize() != 32)
            throw std::runtime_error("Vocabulary::transform features are of different size than the vocabulary ones");
        const int n = features.rows;
        std::vector<uint8_t> desc((size_t)n * 32);
        for (int i = 0; i < n; i++) std::memcpy(&desc[(size_t)i * 32], features.ptr(i), 32);
        std::vector<uint32_t> wid(n), nid(n), feat(n);
        std::vector<float> ww(n);
        std::vector<int> noff(n + 1);
        int nw = 0, nn = 0;
        orbv_set_memory_space(h_, ORBX_MEM_HOST, ORBX_MEM_HOST);
        if (orbv_transform_batch(h_, desc.data(), &n, 1, n, level, normalize, wid.data(), ww.data(), &nw, nid.data(),
                                 noff.data(), feat.data(), &nn) != ORBX_OK)
            throw std::runtime_error(std::string("Vocabulary::transform ") + orbx_last_error_string());
        r1.clear();
        for (int i = 0; i < nw; i++) r1[wid[i]].var = ww[i];
        if (r2) {
            r2->clear();
            for (int j = 0; j < nn; j++) (*r2)[nid[j]].assign(feat.begin() + noff[j], feat.begin() + noff[j + 1]);
        }
    }
    orbv_t* h_ = nullptr;
    int device_ = 0;
};

}  // namespace fbow
#endif
