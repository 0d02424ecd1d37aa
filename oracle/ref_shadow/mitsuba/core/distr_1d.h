// TEST INFRASTRUCTURE (oracle/): wraps the reference's <mitsuba/core/distr_1d.h>.
//
// The real header is included unchanged (its DiscreteDistribution is renamed on the fly) and is
// what scalar instantiations (Float = float, i.e. scalar_rgb / scalar_spectral) use. For array
// instantiations (Float = dr::DynamicArray<float>, standing in for llvm_ad_*) the reference would
// take its `dr::is_jit_v` branches, which cannot be instantiated without the JIT. Those branches
// are restated here per lane:
//   * compute_cdf        (distr_1d.h:218-231): fp32 inclusive prefix sum, sum = cdf[n-1],
//                         normalization = rcp(sum) = 1/sum on the LLVM backend
//                         (drjit-core llvm_red.h:128-168 runs it sequentially for n = 20)
//   * sample             (distr_1d.h:116-134): binary search over [0, n-1] with the JIT predicate
//                         ((cdf < u*sum) || cdf == 0) && cdf != sum   (drjit/util.h:136-181)
//   * sample_reuse       (distr_1d.h:172-182): (u - cdf[i-1]*norm) / (pmf[i]*norm)
#pragma once
#include <drjit/util.h>
#include <drjit/dynamic.h>
// ContinuousDistribution::sample_pdf (distr_1d.h:473) hands its `m_valid` bounds to
// dr::binary_search as width-1 arrays; JIT types have an overload for that, DynamicArray does not.
namespace drjit {
template <typename IndexN, typename Predicate>
IndexN binary_search(const DynamicArray<uint32_t> &start, const DynamicArray<uint32_t> &end, const Predicate &pred) {
    return binary_search<IndexN>(start[0], end[0], pred);
}
}
#define DiscreteDistribution DiscreteDistributionRef
#include_next <mitsuba/core/distr_1d.h>
#undef DiscreteDistribution

NAMESPACE_BEGIN(mitsuba)

template <typename Value> struct DiscreteDistributionJitFlavour {
    using Float = Value;
    using FloatStorage = DynamicBuffer<Float>;
    using Index = dr::uint32_array_t<Value>;
    using Mask = dr::mask_t<Value>;

    DiscreteDistributionJitFlavour() { }
    DiscreteDistributionJitFlavour(const FloatStorage &pmf) : m_pmf(pmf) {
        size_t n = pmf.size();
        if (n == 0) Throw("DiscreteDistribution: empty distribution!");
        m_cdf = dr::zeros<FloatStorage>(n);
        float accum = 0.f; bool any_pos = false;
        for (size_t i = 0; i < n; ++i) {
            float v = pmf[i];
            if (!(v >= 0.f)) Throw("DiscreteDistribution: entries must be non-negative!");
            any_pos |= v > 0.f;
            accum = accum + v;
            m_cdf[i] = accum;
        }
        if (!any_pos) Throw("DiscreteDistribution: no probability mass found!");
        m_sum = m_cdf[n - 1];
        m_normalization = 1.f / m_sum;
    }

    const FloatStorage &pmf() const { return m_pmf; }
    const FloatStorage &cdf() const { return m_cdf; }
    float sum() const { return m_sum; }
    float normalization() const { return m_normalization; }
    size_t size() const { return m_pmf.size(); }

    uint32_t sample_lane(float u) const {
        float s = u * m_sum;
        uint32_t start = 0, end = (uint32_t) m_pmf.size() - 1;
        uint32_t iterations = (start < end) ? (dr::log2i(end - start) + 1) : 0;
        for (uint32_t it = 0; it < iterations; ++it) {
            uint32_t middle = (start + end) >> 1;
            float value = m_cdf[middle];
            bool cond = ((value < s) || value == 0.f) && value != m_sum;
            if (cond) start = std::min(middle + 1, end); else end = middle;
        }
        return start;
    }

    std::pair<Index, Value> sample_reuse(Value value, Mask /*active*/ = true) const {
        size_t n = value.size();
        Index idx = dr::zeros<Index>(n);
        Value out = dr::zeros<Value>(n);
        for (size_t l = 0; l < n; ++l) {
            float u = value[l];
            uint32_t i = sample_lane(u);
            float pmf = m_pmf[i] * m_normalization,
                  cdf = i > 0 ? m_cdf[i - 1] * m_normalization : 0.f;
            idx[l] = i;
            out[l] = (u - cdf) / pmf;
        }
        return { idx, out };
    }

private:
    FloatStorage m_pmf, m_cdf;
    float m_sum = 0.f, m_normalization = 0.f;
};

namespace detail {
    template <typename Value, bool IsArray = dr::is_array_v<Value>> struct discrete_distr_sel {
        using type = DiscreteDistributionRef<Value>;
    };
    template <typename Value> struct discrete_distr_sel<Value, true> {
        using type = DiscreteDistributionJitFlavour<Value>;
    };
}
template <typename Value> struct DiscreteDistribution : detail::discrete_distr_sel<Value>::type {
    using Base = typename detail::discrete_distr_sel<Value>::type;
    using Base::Base;
    DiscreteDistribution() { }
};

NAMESPACE_END(mitsuba)
