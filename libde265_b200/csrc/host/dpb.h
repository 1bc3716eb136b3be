// dpb.h — reference picture bookkeeping on the host: reference picture set (H.265 8.3.2), picture marking and
// reference picture list construction (8.3.4).  Counterpart of the reference's decctx.cc:1090-1500
// (process_reference_picture_set, construct_reference_picture_lists).  The pictures themselves live in HBM
// (RefPic::surface); this is only the bookkeeping, shared by the decoder and the test-stream generator.
#pragma once
#include <functional>
#include <memory>
#include <vector>
#include "picture.h"

namespace b265h {

struct Dpb {
  std::vector<std::shared_ptr<RefPic>> pics;   // pictures marked "used for reference"
  // reference picture set of the current picture
  std::vector<std::shared_ptr<RefPic>> st_before, st_after, lt_curr;
  bool missing = false;                        // a picture of the set is not in the buffer (and could not be substituted)
  // Substitute for a picture of the set that is not in the buffer (8.3.3.2, decctx.cc:1408-1434: a mid-grey picture
  // whose motion field reads "all intra", never output).  Unset: such a picture makes the lists fail.
  std::function<std::shared_ptr<RefPic>(int poc, bool long_term)> make_unavailable;
  std::shared_ptr<RefPic> substitute(int poc, bool lt) {
    if (!make_unavailable) { missing = true; return nullptr; }
    std::shared_ptr<RefPic> r = make_unavailable(poc, lt);
    if (!r) missing = true;
    return r;
  }

  void clear() { pics.clear(); st_before.clear(); st_after.clear(); lt_curr.clear(); missing = false; }

  std::shared_ptr<RefPic> find_st(int poc) const {
    for (const auto& p : pics) if (!p->is_lt && p->poc == poc) return p;
    return nullptr;
  }
  std::shared_ptr<RefPic> find_any(int poc, int mask) const {
    for (const auto& p : pics) if ((p->poc & mask) == poc) return p;
    return nullptr;
  }

  // 8.3.2, once per picture (from its first slice segment header)
  void apply_rps(const SliceHeader& sh, const Sps& sps, int poc, bool idr_like) {
    st_before.clear(); st_after.clear(); lt_curr.clear();
    missing = false;
    if (idr_like) { pics.clear(); return; }
    const bool idr = sh.nal_type == NAL_IDR_W_RADL || sh.nal_type == NAL_IDR_N_LP;
    if (idr) { pics.clear(); return; }
    std::vector<std::shared_ptr<RefPic>> keep;
    const int max_lsb = 1 << sps.log2_max_poc_lsb;
    // long-term pictures first: they are identified among ALL reference pictures and then re-marked
    for (int i = 0; i < sh.num_long_term_sps + sh.num_long_term_pics; i++) {
      int p = sh.poc_lsb_lt[i];
      std::shared_ptr<RefPic> r;
      if (sh.delta_poc_msb_present[i]) {
        p += poc - sh.delta_poc_msb_cycle[i] * max_lsb - (poc & (max_lsb - 1));
        r = find_any(p, ~0);
      } else {
        r = find_any(p, max_lsb - 1);
      }
      // a missing long-term picture is generated whether the current picture uses it or not (decctx.cc:1572-1612),
      // with the POC the header gives (LSB only when the MSB is not signalled)
      if (!r) r = substitute(p, true);
      if (r) { r->is_lt = true; keep.push_back(r); }
      if (sh.used_by_curr_lt[i]) lt_curr.push_back(r);
    }
    const ShortTermRps& rps = sh.st_rps;
    for (int i = 0; i < rps.num_negative; i++) {
      std::shared_ptr<RefPic> r = find_st(poc + rps.delta_poc_s0[i]);
      if (!r && rps.used_s0[i]) r = substitute(poc + rps.delta_poc_s0[i], false);   // only pictures of the Curr sets (decctx.cc:1627-1667)
      if (r) keep.push_back(r);
      if (rps.used_s0[i]) st_before.push_back(r);
    }
    for (int i = 0; i < rps.num_positive; i++) {
      std::shared_ptr<RefPic> r = find_st(poc + rps.delta_poc_s1[i]);
      if (!r && rps.used_s1[i]) r = substitute(poc + rps.delta_poc_s1[i], false);
      if (r) keep.push_back(r);
      if (rps.used_s1[i]) st_after.push_back(r);
    }
    pics.swap(keep);   // everything else is no longer used for reference
  }

  // 8.3.4 for one slice; fills pic.ref_pics (slot table) and returns the slice's lists. false: a picture is missing.
  bool build_lists(const SliceHeader& sh, Picture& pic, SliceRefs& out) const {
    out = SliceRefs();
    if (sh.slice_type == SLICE_I) return true;
    const int total = (int)(st_before.size() + st_after.size() + lt_curr.size());
    if (total == 0 || missing) return false;
    for (int X = 0; X < (sh.slice_type == SLICE_B ? 2 : 1); X++) {
      std::vector<const std::shared_ptr<RefPic>*> temp;
      std::vector<uint8_t> temp_lt;
      const int n_active = sh.num_ref_idx[X];
      const int n_temp = std::max(n_active, total);
      const auto& first = X ? st_after : st_before;
      const auto& second = X ? st_before : st_after;
      while ((int)temp.size() < n_temp) {
        for (const auto& r : first) if ((int)temp.size() < n_temp) { temp.push_back(&r); temp_lt.push_back(0); }
        for (const auto& r : second) if ((int)temp.size() < n_temp) { temp.push_back(&r); temp_lt.push_back(0); }
        for (const auto& r : lt_curr) if ((int)temp.size() < n_temp) { temp.push_back(&r); temp_lt.push_back(1); }
      }
      out.n[X] = n_active;
      for (int i = 0; i < n_active; i++) {
        const int e = sh.ref_pic_list_modification[X] ? sh.list_entry[X][i] : i;
        if (e >= n_temp) return false;
        const std::shared_ptr<RefPic>& r = *temp[e];
        if (!r) return false;
        out.poc[X][i] = r->poc;
        out.is_lt[X][i] = temp_lt[e];
        size_t slot = 0;
        while (slot < pic.ref_pics.size() && pic.ref_pics[slot] != r) slot++;
        if (slot == pic.ref_pics.size()) {
          if (slot >= 16) return false;   // the back end holds 16 reference slots per picture (FrameDev::ref_plane)
          pic.ref_pics.push_back(r);
        }
        out.slot[X][i] = (uint8_t)slot;
      }
    }
    return true;
  }
};

}  // namespace b265h
