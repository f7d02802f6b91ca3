// Bounding-box COCO evaluation (the 12 summary numbers of pycocotools.COCOeval(iouType='bbox')).
//
// Behavioural spec (public COCO evaluation protocol as implemented by pycocotools, which the
// reference pulls in at train.py:135 through keras_retinanet.callbacks.coco.CocoEval; SURVEY.md
// B18, N7): per image and category, detections sorted by score (at most 100), greedy matching at
// IoU thresholds .50:.05:.95 with crowd regions matchable many times (IoU against a crowd uses the
// detection's own area as the union), ground truth outside the area range or marked crowd is
// "ignore"; precision is made monotone and sampled at 101 recall points; averages skip entries with
// no ground truth (-1).  Only boxes are handled (no mask API is needed for bbox evaluation).
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <map>
#include <numeric>
#include <vector>

#include "../../include/b200_rt.h"

namespace {

struct Box {
  double x, y, w, h, area, score;
  int crowd, ignore;
  long long id;
};

struct ImgEval {
  // per IoU threshold: detection matched flag / ignore flag; shared: scores, gt ignore flags
  std::vector<double> scores;
  std::vector<std::vector<uint8_t>> dt_matched, dt_ignore;  // [T][D]
  int num_gt_not_ignored = 0;
};

constexpr int T = 10, R = 101, A = 4, M = 3;
const double kArea[A][2] = {{0, 1e10}, {0, 32.0 * 32}, {32.0 * 32, 96.0 * 96}, {96.0 * 96, 1e10}};
const int kMaxDets[M] = {1, 10, 100};

double iou(const Box& d, const Box& g) {
  const double iw = std::min(d.x + d.w, g.x + g.w) - std::max(d.x, g.x);
  const double ih = std::min(d.y + d.h, g.y + g.h) - std::max(d.y, g.y);
  if (iw <= 0 || ih <= 0) return 0.0;
  const double inter = iw * ih;
  const double uni = g.crowd ? d.w * d.h : d.w * d.h + g.w * g.h - inter;
  return uni > 0 ? inter / uni : 0.0;
}

ImgEval evaluate_img(const std::vector<Box>& gts_in, const std::vector<Box>& dts_sorted, const double* rng, int max_det) {
  ImgEval e;
  std::vector<Box> gts = gts_in;
  for (Box& g : gts) g.ignore = (g.ignore || g.crowd || g.area < rng[0] || g.area > rng[1]) ? 1 : 0;
  std::stable_sort(gts.begin(), gts.end(), [](const Box& a, const Box& b) { return a.ignore < b.ignore; });
  const int D = std::min<int>(max_det, static_cast<int>(dts_sorted.size()));
  const int G = static_cast<int>(gts.size());
  std::vector<double> ious(static_cast<size_t>(D) * G);
  for (int d = 0; d < D; ++d)
    for (int g = 0; g < G; ++g) ious[static_cast<size_t>(d) * G + g] = iou(dts_sorted[d], gts[g]);
  e.dt_matched.assign(T, std::vector<uint8_t>(D, 0));
  e.dt_ignore.assign(T, std::vector<uint8_t>(D, 0));
  e.scores.resize(D);
  for (int d = 0; d < D; ++d) e.scores[d] = dts_sorted[d].score;
  for (const Box& g : gts) e.num_gt_not_ignored += !g.ignore;
  for (int t = 0; t < T; ++t) {
    const double thr = 0.5 + 0.05 * t;
    std::vector<uint8_t> gt_matched(G, 0);
    for (int d = 0; d < D; ++d) {
      double best = std::min(thr, 1 - 1e-10);
      int m = -1;
      for (int g = 0; g < G; ++g) {
        if (gt_matched[g] && !gts[g].crowd) continue;
        if (m > -1 && !gts[m].ignore && gts[g].ignore) break;  // only worse (ignored) candidates remain
        const double v = ious[static_cast<size_t>(d) * G + g];
        if (v < best) continue;
        best = v;
        m = g;
      }
      if (m == -1) continue;
      e.dt_ignore[t][d] = gts[m].ignore;
      e.dt_matched[t][d] = 1;
      gt_matched[m] = 1;
    }
    for (int d = 0; d < D; ++d) {
      const double a = dts_sorted[d].w * dts_sorted[d].h;
      if (!e.dt_matched[t][d] && (a < rng[0] || a > rng[1])) e.dt_ignore[t][d] = 1;
    }
  }
  return e;
}

double mean_valid(const std::vector<double>& v) {
  double s = 0;
  long long n = 0;
  for (double x : v)
    if (x > -1) {
      s += x;
      ++n;
    }
  return n ? s / n : -1.0;
}

}  // namespace

extern "C" int b200rt_cocoeval_bbox(const double* gt, const double* gt_area, long long ng, const double* dt,
                                    long long nd, const long long* img_ids, long long n_img, const long long* cat_ids,
                                    long long n_cat, double* stats_out) {
  std::map<std::pair<long long, long long>, std::vector<Box>> G, Dm;
  for (long long i = 0; i < ng; ++i) {
    const double* r = gt + 7 * i;
    Box b{r[2], r[3], r[4], r[5], gt_area ? gt_area[i] : r[4] * r[5], 0.0, (static_cast<int>(r[6]) & 1),
          (static_cast<int>(r[6]) >> 1) & 1, i + 1};
    G[{static_cast<long long>(r[0]), static_cast<long long>(r[1])}].push_back(b);
  }
  for (long long i = 0; i < nd; ++i) {
    const double* r = dt + 7 * i;
    Box b{r[2], r[3], r[4], r[5], r[4] * r[5], r[6], 0, 0, i + 1};
    Dm[{static_cast<long long>(r[0]), static_cast<long long>(r[1])}].push_back(b);
  }
  for (auto& kv : Dm)
    std::stable_sort(kv.second.begin(), kv.second.end(), [](const Box& a, const Box& b) { return a.score > b.score; });

  // precision[T][R][K][A][M], recall[T][K][A][M]
  const size_t K = static_cast<size_t>(n_cat);
  std::vector<double> precision(static_cast<size_t>(T) * R * K * A * M, -1.0), recall(static_cast<size_t>(T) * K * A * M, -1.0);
  auto P = [&](int t, int r, size_t k, int a, int m) -> double& { return precision[(((static_cast<size_t>(t) * R + r) * K + k) * A + a) * M + m]; };
  auto Rc = [&](int t, size_t k, int a, int m) -> double& { return recall[((static_cast<size_t>(t) * K + k) * A + a) * M + m]; };
  static const std::vector<Box> kEmpty;
  for (size_t k = 0; k < K; ++k) {
    for (int a = 0; a < A; ++a) {
      // evaluate every image once at maxDet = 100, then truncate per maxDet when accumulating
      std::vector<ImgEval> evals;
      for (long long ii = 0; ii < n_img; ++ii) {
        auto key = std::make_pair(img_ids[ii], cat_ids[k]);
        auto gi = G.find(key);
        auto di = Dm.find(key);
        if (gi == G.end() && di == Dm.end()) continue;
        evals.push_back(evaluate_img(gi == G.end() ? kEmpty : gi->second, di == Dm.end() ? kEmpty : di->second,
                                     kArea[a], kMaxDets[M - 1]));
      }
      for (int m = 0; m < M; ++m) {
        const int md = kMaxDets[m];
        int npig = 0;
        struct Det { double score; int e, d; };
        std::vector<Det> dets;
        for (size_t e = 0; e < evals.size(); ++e) {
          npig += evals[e].num_gt_not_ignored;
          const int D = std::min<int>(md, static_cast<int>(evals[e].scores.size()));
          for (int d = 0; d < D; ++d) dets.push_back({evals[e].scores[d], static_cast<int>(e), d});
        }
        if (npig == 0) continue;
        std::stable_sort(dets.begin(), dets.end(), [](const Det& x, const Det& y) { return x.score > y.score; });
        for (int t = 0; t < T; ++t) {
          std::vector<double> rc, pr;
          double tp = 0, fp = 0;
          for (const Det& q : dets) {
            const ImgEval& e = evals[q.e];
            if (e.dt_ignore[t][q.d]) continue;  // ignored detections do not count either way
            if (e.dt_matched[t][q.d]) tp += 1; else fp += 1;
            rc.push_back(tp / npig);
            pr.push_back(tp / (fp + tp + 2.220446049250313e-16));
          }
          const int nd2 = static_cast<int>(rc.size());
          Rc(t, k, a, m) = nd2 ? rc.back() : 0.0;
          for (int i = nd2 - 1; i > 0; --i)
            if (pr[i] > pr[i - 1]) pr[i - 1] = pr[i];
          for (int r = 0; r < R; ++r) {
            const double thr = r * 0.01;
            const int pi = static_cast<int>(std::lower_bound(rc.begin(), rc.end(), thr) - rc.begin());
            P(t, r, k, a, m) = pi < nd2 ? pr[pi] : 0.0;
          }
        }
      }
    }
  }
  auto ap = [&](int t_lo, int t_hi, int a, int m) {
    std::vector<double> v;
    for (int t = t_lo; t < t_hi; ++t)
      for (int r = 0; r < R; ++r)
        for (size_t k = 0; k < K; ++k) v.push_back(P(t, r, k, a, m));
    return mean_valid(v);
  };
  auto ar = [&](int a, int m) {
    std::vector<double> v;
    for (int t = 0; t < T; ++t)
      for (size_t k = 0; k < K; ++k) v.push_back(Rc(t, k, a, m));
    return mean_valid(v);
  };
  stats_out[0] = ap(0, T, 0, 2);
  stats_out[1] = ap(0, 1, 0, 2);
  stats_out[2] = ap(5, 6, 0, 2);
  stats_out[3] = ap(0, T, 1, 2);
  stats_out[4] = ap(0, T, 2, 2);
  stats_out[5] = ap(0, T, 3, 2);
  stats_out[6] = ar(0, 0);
  stats_out[7] = ar(0, 1);
  stats_out[8] = ar(0, 2);
  stats_out[9] = ar(1, 2);
  stats_out[10] = ar(2, 2);
  stats_out[11] = ar(3, 2);
  return 0;
}
