// TEST INFRASTRUCTURE (oracle/): stand-in for <mitsuba/core/properties.h>: a typed name->value bag
// with the accessors SunskyEmitter::init_from_props uses (reference: src/emitters/sunsky.cpp:889-948).
#pragma once
#include <mitsuba/core/object.h>
#include <mitsuba/core/logger.h>
#include <mitsuba/core/vector.h>
#include <mitsuba/core/transform.h>
#include <map>
#include <string>
#include <array>

NAMESPACE_BEGIN(mitsuba)

class Properties {
public:
    void set_float(const std::string &k, double v) { m_float[k] = v; }
    void set_vec3(const std::string &k, float x, float y, float z) { m_vec3[k] = { x, y, z }; }
    /// The real Properties stores transforms as Transform<Point<double, 4>> (`using Float = double`,
    /// reference include/mitsuba/core/properties.h:70, 54): the inverse-transpose a ScalarTransform4f carries is
    /// the fp64 one (dr::inverse_transpose on doubles, transform.h:55-57) rounded to fp32 by the converting
    /// constructor (transform.h:64-67). `inv16` (nullable): an inverse the caller built itself, e.g. a Python
    /// Transform4f whose factory filled both matrices analytically (transform.h:59-62).
    void set_matrix(const std::string &k, const float *m16, const float *inv16 = nullptr) {
        using TransformD = Transform<Point<double, 4>>;
        typename TransformD::Matrix m, it;
        for (int r = 0; r < 4; ++r)
            for (int c = 0; c < 4; ++c) {
                m(r, c) = (double) m16[r * 4 + c];
                if (inv16) it(c, r) = (double) inv16[r * 4 + c];      // stored transposed
            }
        m_mat.erase(k);
        m_mat.emplace(k, inv16 ? TransformD(m, it) : TransformD(m));
    }
    void set_albedo(const float *v, int n) { m_albedo.assign(v, v + n); }
    const std::vector<float> &albedo() const { return m_albedo; }

    bool has_property(const std::string &k) const {
        return m_float.count(k) || m_vec3.count(k) || m_mat.count(k);
    }

    template <typename T> T get(const std::string &k) const {
        if constexpr (std::is_arithmetic_v<T>) {
            auto it = m_float.find(k);
            if (it == m_float.end()) Throw("Property \"%s\" has not been specified!", k);
            return (T) it->second;
        } else if constexpr (dr::size_v<T> == 3 && dr::depth_v<T> == 1) {
            auto it = m_vec3.find(k);
            if (it == m_vec3.end()) Throw("Property \"%s\" has not been specified!", k);
            return T(it->second[0], it->second[1], it->second[2]);
        } else {
            // 4x4 affine transform: fp64 matrix and inverse-transpose, both cast (src/core/properties.cpp:129-131)
            auto it = m_mat.find(k);
            if (it == m_mat.end()) Throw("Property \"%s\" has not been specified!", k);
            return T(it->second);
        }
    }

    template <typename T, typename D> T get(const std::string &k, const D &def) const {
        if (!has_property(k)) {
            if constexpr (std::is_arithmetic_v<T>) return (T) def; else return T(def);
        }
        return get<T>(k);
    }

    /// The albedo is always handed over as per-channel values (3 or 11); see render/texture.h
    template <typename Texture, typename D> ref<Texture> texture(const std::string &, const D &def) const {
        if (m_albedo.empty())
            return new Texture(std::vector<float>(1, (float) def));
        return new Texture(m_albedo);
    }

private:
    std::map<std::string, double> m_float;
    std::map<std::string, std::array<float, 3>> m_vec3;
    std::map<std::string, Transform<Point<double, 4>>> m_mat;
    std::vector<float> m_albedo;
};

NAMESPACE_END(mitsuba)
