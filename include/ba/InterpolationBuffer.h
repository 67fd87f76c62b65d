/*
 * ba/InterpolationBuffer.h — time-keyed buffer that feeds AddImuResidual.
 *
 * Same interface and results as ba::InterpolationBufferT of the reference
 * (reference include/ba/InterpolationBuffer.h:38-210): public fields elements / start_time / end_time /
 * average_dt, AddElement, Clear, HasElement, GetElement (two overloads), GetNext, GetRange.  The caller
 * pattern it serves is imu_buffer.GetRange(pose_i.time, pose_j.time) -> AddImuResidual(i, j, measurements)
 * (reference applications/unary_binary_imu_test/main.cpp:243-280).
 *
 * Inside it differs: the reference guesses an index from average_dt and walks element by element to the
 * bracketing pair; here the bracket is found by binary search on the (strictly increasing) time key, on
 * the side of the same guess, so every call is O(log n) however irregular the sampling is, and the
 * returned index is the one the reference's walk ends on — including when `time` equals a stored
 * time stamp, where the reference's answer depends on which side its guess falls (the interpolated value
 * is the stored element either way; the index, and with it whether GetRange repeats that element, is not).
 * Two cases the reference leaves undefined are defined here: a query before the first element returns
 * the first element, a query at or after the last returns the last (the reference reads one past the
 * end when its guess is below the last index and time > end_time).
 *
 * New for the GPU path: AppendRangeFlat writes a range as the 7-double rows (wx wy wz ax ay az time) that
 * bagpu_set_imu_residuals takes (include/ba_b200.h), without an intermediate vector of elements.
 *
 * ElementType needs: a `time` member, operator*(ScalarType) and operator+(ElementType).
 */
#ifndef BA_B200_INTERPOLATIONBUFFER_H
#define BA_B200_INTERPOLATIONBUFFER_H

#include <algorithm>
#include <cassert>
#include <cstddef>
#include <vector>

namespace ba {

template <typename ElementType, typename ScalarType>
struct InterpolationBufferT {
  std::vector<ElementType> elements;
  ScalarType start_time;
  ScalarType end_time;
  ScalarType average_dt;

  InterpolationBufferT(unsigned int size = 1000) : start_time(-1), end_time(-1), average_dt(-1) { elements.reserve(size); }

  void Clear() {
    elements.clear();
    start_time = end_time = average_dt = -1;
  }

  /* reference InterpolationBuffer.h:64-77.  average_dt is the running mean over n+1 samples in which the
   * first element counts with dt = 0 — kept, because the index guess below depends on it. */
  void AddElement(const ElementType& element) {
    assert(element.time > end_time);
    const size_t n = elements.size();
    const ScalarType dt = n ? element.time - elements.back().time : ScalarType(0);
    average_dt = (average_dt == -1) ? dt : (average_dt * n + dt) / (n + 1);
    elements.push_back(element);
    start_time = elements.front().time;
    end_time = element.time;
  }

  bool HasElement(const ScalarType time) const { return time >= start_time && time <= end_time; }

  ElementType GetElement(const ScalarType time) const {
    size_t index;
    return GetElement(time, &index);
  }

  /* reference InterpolationBuffer.h:134-182 */
  ElementType GetElement(const ScalarType time, size_t* index_out) const {
    const size_t n = elements.size();
    assert(n > 0);
    if (n == 1 || time < elements.front().time) { *index_out = 0; return elements.front(); }
    if (time >= elements.back().time) { *index_out = n - 1; return elements.back(); }
    /* from here: front.time <= time < back.time, so a bracketing pair exists */
    const size_t guess = GuessIndex(time);
    size_t lo;   /* the pair is (lo, lo + 1) */
    if (elements[guess].time > time) {
      /* the reference walks down to the first pair whose left time is <= time */
      lo = UpperBound(0, guess, time) - 1;
    } else {
      /* the reference walks up while the next time is still < time: it stays on `guess` if that already
       * brackets, which is what makes an exact hit on elements[guess].time come back with index guess */
      const size_t below = LowerBound(guess, n, time);   /* first index in [guess, n) with time stamp >= time */
      lo = below > guess ? below - 1 : guess;
    }
    *index_out = lo;
    const ElementType& a = elements[lo];
    const ElementType& b = elements[lo + 1];
    const ScalarType s = (time - a.time) / (b.time - a.time);
    ElementType res = a * (1 - s) + b * s;
    res.time = time;
    return res;
  }

  /* reference InterpolationBuffer.h:89-104: the next stored element if its time is <= max_time (true), else
   * the element interpolated at max_time (false = the range is finished) */
  bool GetNext(const ScalarType max_time, size_t& index, ElementType& output) const {
    if (index + 1 < elements.size() && !(elements[index + 1].time > max_time)) {
      output = elements[++index];
      return true;
    }
    output = GetElement(max_time, &index);
    return false;
  }

  /* reference InterpolationBuffer.h:184-210: [start, end] clipped to the buffer; first and last entries are
   * interpolated at the clipped bounds, the stored elements in between are returned as they are */
  std::vector<ElementType> GetRange(ScalarType start, ScalarType end) const {
    std::vector<ElementType> out;
    start = std::max(start, start_time);
    end = std::min(end, end_time);
    if (!HasElement(start)) return out;
    size_t index;
    out.push_back(GetElement(start, &index));
    ElementType e;
    bool more;
    do {
      more = GetNext(end, index, e);
      out.push_back(e);
    } while (more);
    return out;
  }

  /* The same range as rows of 7 doubles (w, a, time) appended to `meas7`: the layout of
   * bagpu_set_imu_residuals (include/ba_b200.h).  Returns the number of rows appended.  ElementType must have
   * 3-vectors `w` and `a` indexable with []. */
  template <typename Vec>
  size_t AppendRangeFlat(ScalarType start, ScalarType end, Vec& meas7) const {
    start = std::max(start, start_time);
    end = std::min(end, end_time);
    if (!HasElement(start)) return 0;
    size_t index, rows = 0;
    auto put = [&](const ElementType& m) {
      for (int k = 0; k < 3; ++k) meas7.push_back((double)m.w[k]);
      for (int k = 0; k < 3; ++k) meas7.push_back((double)m.a[k]);
      meas7.push_back((double)m.time);
      ++rows;
    };
    put(GetElement(start, &index));
    ElementType e;
    bool more;
    do {
      more = GetNext(end, index, e);
      put(e);
    } while (more);
    return rows;
  }

 private:
  /* reference InterpolationBuffer.h:137-140: (time - start_time) / average_dt, truncated and clipped */
  size_t GuessIndex(const ScalarType time) const {
    const size_t n = elements.size();
    const ScalarType g = (time - start_time) / average_dt;
    if (!(g > 0)) return 0;
    return (g >= ScalarType(n - 1)) ? n - 1 : (size_t)g;
  }
  /* first index in [lo, hi) whose time stamp is > time (hi if none) */
  size_t UpperBound(size_t lo, size_t hi, const ScalarType time) const {
    while (lo < hi) { const size_t mid = lo + (hi - lo) / 2; if (elements[mid].time > time) hi = mid; else lo = mid + 1; }
    return lo;
  }
  /* first index in [lo, hi) whose time stamp is >= time (hi if none) */
  size_t LowerBound(size_t lo, size_t hi, const ScalarType time) const {
    while (lo < hi) { const size_t mid = lo + (hi - lo) / 2; if (elements[mid].time < time) lo = mid + 1; else hi = mid; }
    return lo;
  }
};

}  // namespace ba

#endif /* BA_B200_INTERPOLATIONBUFFER_H */
