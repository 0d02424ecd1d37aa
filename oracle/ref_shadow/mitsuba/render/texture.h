// TEST INFRASTRUCTURE (oracle/): stand-in for <mitsuba/render/texture.h>: a spatially uniform
// albedo with one value per sunsky channel (3 RGB / 11 spectral bins at 320..720 nm), or a single
// value for all of them. Only eval() and is_spatially_varying() are used
// (reference: src/emitters/sunsky.cpp:956-978).
#pragma once
#include <mitsuba/core/object.h>
#include <mitsuba/core/spectrum.h>
#include <mitsuba/render/interaction.h>
#include <vector>

NAMESPACE_BEGIN(mitsuba)

template <typename Float, typename Spectrum> class Texture : public Object {
public:
    MI_IMPORT_TYPES()
    Texture(const std::vector<float> &values) : m_values(values) { }
    bool is_spatially_varying() const { return false; }

    UnpolarizedSpectrum eval(const SurfaceInteraction3f &si, Mask = true) const {
        if constexpr (is_rgb_v<Spectrum>) {
            if (m_values.size() == 1) return UnpolarizedSpectrum(m_values[0]);
            return UnpolarizedSpectrum(m_values[0], m_values[1], m_values[2]);
        } else {
            if (m_values.size() == 1) return UnpolarizedSpectrum(m_values[0]);
            // nearest bin of the 40 nm grid starting at 320 nm
            UnpolarizedSpectrum r;
            for (size_t c = 0; c < r.size(); ++c) {
                if constexpr (dr::is_jit_v<Float>) {            // JIT arrays (adapter type-check): lanes through a host copy
                    size_t n = si.wavelengths[c].size();
                    std::vector<float> tmp(n);
                    for (size_t l = 0; l < n; ++l) {
                        int b = (int) ((si.wavelengths[c].entry(l) - 320.f) / 40.f + 0.5f);
                        tmp[l] = m_values[std::min(std::max(b, 0), (int) m_values.size() - 1)];
                    }
                    r[c] = dr::load<Float>(tmp.data(), n);
                } else if constexpr (dr::is_array_v<Float>) {
                    size_t n = si.wavelengths[c].size();
                    Float v = dr::zeros<Float>(n);
                    for (size_t l = 0; l < n; ++l) {
                        int b = (int) ((si.wavelengths[c][l] - 320.f) / 40.f + 0.5f);
                        v[l] = m_values[std::min(std::max(b, 0), (int) m_values.size() - 1)];
                    }
                    r[c] = v;
                } else {
                    int b = (int) ((si.wavelengths[c] - 320.f) / 40.f + 0.5f);
                    r[c] = m_values[std::min(std::max(b, 0), (int) m_values.size() - 1)];
                }
            }
            return r;
        }
    }
private:
    std::vector<float> m_values;
};

NAMESPACE_END(mitsuba)
