// bow_ref.cc — see bow_ref.h (TEST INFRASTRUCTURE: only tests/, smoke() and bench.py's CPU legs use it).
#include "bow_ref.h"

#include <cmath>
#include <map>

#include "match_ref.h"

namespace orbref {

bool VocabularyRef::build(int k_, int L_, int weighting_, int norm_, int nnodes, const int32_t* par, const uint8_t* isLeaf,
                          const uint8_t* d, const double* w) {
    if (nnodes < 1 || k_ < 1 || L_ < 1) return false;
    k = k_; L = L_; weighting = weighting_; norm = norm_;
    parent.assign(par, par + nnodes);
    desc.assign(d, d + 32 * (size_t)nnodes);
    weight.assign(w, w + nnodes);
    childCount.assign(nnodes, 0);
    childBase.assign(nnodes + 1, 0);
    wordId.assign(nnodes, -1);
    for (int i = 1; i < nnodes; ++i) {
        if (par[i] < 0 || par[i] >= i) return false;
        childCount[par[i]]++;
    }
    for (int i = 0; i < nnodes; ++i) childBase[i + 1] = childBase[i] + childCount[i];
    childIdx.assign(childBase[nnodes], 0);
    std::vector<int32_t> fill(childBase.begin(), childBase.end() - 1);
    for (int i = 1; i < nnodes; ++i) childIdx[fill[par[i]]++] = i;          // children.push_back(nid), file order
    nwords = 0;
    for (int i = 1; i < nnodes; ++i) {
        if ((isLeaf[i] != 0) != (childCount[i] == 0)) return false;          // flag must agree with isLeaf() = children.empty()
        if (isLeaf[i]) wordId[i] = nwords++;
    }
    return childCount[0] > 0;
}

void VocabularyRef::transformOne(const uint8_t* feature, int levelsup, uint32_t& word, double& w, uint32_t& nid) const {
    const int nidLevel = L - levelsup;
    nid = 0;
    int finalId = 0, level = 0;
    do {
        ++level;
        const int b = childBase[finalId], cnt = childCount[finalId];
        finalId = childIdx[b];
        double bestD = descriptorDistanceSwar(feature, &desc[32 * (size_t)finalId]);
        for (int j = 1; j < cnt; ++j) {
            const int id = childIdx[b + j];
            const double dd = descriptorDistanceSwar(feature, &desc[32 * (size_t)id]);
            if (dd < bestD) { bestD = dd; finalId = id; }
        }
        if (level == nidLevel) nid = (uint32_t)finalId;
    } while (childCount[finalId] != 0);
    if (nidLevel > 0 && level < nidLevel) nid = (uint32_t)finalId;   // defined deviation, see header
    word = (uint32_t)wordId[finalId];
    w = weight[finalId];
}

void bowTransform(const VocabularyRef& v, const uint8_t* desc, int n, int levelsup, std::vector<uint32_t>& words,
                  std::vector<double>& values, std::vector<uint32_t>& nodeIds, std::vector<int32_t>& offsets,
                  std::vector<int32_t>& indices) {
    std::map<uint32_t, double> bow;
    std::map<uint32_t, std::vector<int32_t>> fv;
    const bool tf = v.weighting == BOW_TF || v.weighting == BOW_TF_IDF;
    for (int i = 0; i < n; ++i) {
        uint32_t id, nid;
        double w;
        v.transformOne(desc + 32 * (size_t)i, levelsup, id, w, nid);
        if (w > 0) {
            auto it = bow.lower_bound(id);
            if (it != bow.end() && !(bow.key_comp()(id, it->first))) {
                if (tf) it->second += w;                   // addWeight; addIfNotExist leaves it
            } else {
                bow.insert(it, std::make_pair(id, w));
            }
            fv[nid].push_back(i);
        }
    }
    if (tf && !bow.empty() && v.norm == BOW_NORM_NONE) {
        const double nd = (double)bow.size();
        for (auto& kv : bow) kv.second /= nd;
    }
    if (v.norm != BOW_NORM_NONE) {                         // BowVector::normalize
        double nrm = 0.0;
        if (v.norm == BOW_NORM_L1) { for (auto& kv : bow) nrm += std::fabs(kv.second); }
        else { for (auto& kv : bow) nrm += kv.second * kv.second; nrm = std::sqrt(nrm); }
        if (nrm > 0.0) for (auto& kv : bow) kv.second /= nrm;
    }
    words.clear(); values.clear(); nodeIds.clear(); offsets.assign(1, 0); indices.clear();
    for (auto& kv : bow) { words.push_back(kv.first); values.push_back(kv.second); }
    for (auto& kv : fv) {
        nodeIds.push_back(kv.first);
        indices.insert(indices.end(), kv.second.begin(), kv.second.end());
        offsets.push_back((int32_t)indices.size());
    }
}

}  // namespace orbref
