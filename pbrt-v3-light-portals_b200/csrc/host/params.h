// Typed parameter lists of the pbrt scene format ("integer maxdepth" [5] ...).
// Mirrors the lookup surface of the reference's ParamSet (src/core/paramset.h:
// FindOneInt/Float/Bool/String/Point3f/Spectrum, FindInt/Float/Point3f...,
// ReportUnused) with the same defaulting rule: FindOne* returns the default
// unless the parameter exists with exactly one value (paramset.cpp:310-330).
#ifndef PG_HOST_PARAMS_H
#define PG_HOST_PARAMS_H

#include <string>
#include <vector>

#include "../common/pg_math.h"
#include "spectrum.h"

namespace pg {

enum class ParamType { Int, Bool, Float, Point2, Vector2, Point3, Vector3, Normal, Spectrum, String, Texture };

struct Param {
    ParamType type;
    std::string name;
    std::vector<int> ints;
    std::vector<uint8_t> bools;
    std::vector<float> floats;  // Float, Point2/3, Vector2/3, Normal (flattened)
    std::vector<RGB> spectra;
    std::vector<std::string> strings;  // String, Texture
    mutable bool lookedUp = false;
};

void hostWarning(const char *fmt, ...);
void hostError(const char *fmt, ...);

class ParamSet {
  public:
    void add(Param p) {
        for (auto &q : items)
            if (q.name == p.name && sameClass(q.type, p.type)) {
                q = std::move(p);
                return;
            }
        items.push_back(std::move(p));
    }
    const Param *find(const std::string &name, ParamType t) const {
        for (auto &q : items)
            if (q.name == name && q.type == t) {
                q.lookedUp = true;
                return &q;
            }
        return nullptr;
    }
    int findOneInt(const std::string &n, int d) const {
        auto p = find(n, ParamType::Int);
        return (p && p->ints.size() == 1) ? p->ints[0] : d;
    }
    bool findOneBool(const std::string &n, bool d) const {
        auto p = find(n, ParamType::Bool);
        return (p && p->bools.size() == 1) ? p->bools[0] != 0 : d;
    }
    float findOneFloat(const std::string &n, float d) const {
        auto p = find(n, ParamType::Float);
        return (p && p->floats.size() == 1) ? p->floats[0] : d;
    }
    std::string findOneString(const std::string &n, const std::string &d) const {
        auto p = find(n, ParamType::String);
        return (p && p->strings.size() == 1) ? p->strings[0] : d;
    }
    std::string findTexture(const std::string &n) const {
        auto p = find(n, ParamType::Texture);
        return (p && p->strings.size() == 1) ? p->strings[0] : std::string();
    }
    V3 findOnePoint3(const std::string &n, V3 d) const {
        auto p = find(n, ParamType::Point3);
        return (p && p->floats.size() == 3) ? v3(p->floats[0], p->floats[1], p->floats[2]) : d;
    }
    V3 findOneVector3(const std::string &n, V3 d) const {
        auto p = find(n, ParamType::Vector3);
        return (p && p->floats.size() == 3) ? v3(p->floats[0], p->floats[1], p->floats[2]) : d;
    }
    RGB findOneSpectrum(const std::string &n, RGB d) const {
        auto p = find(n, ParamType::Spectrum);
        return (p && p->spectra.size() == 1) ? p->spectra[0] : d;
    }
    const std::vector<int> *findInts(const std::string &n) const {
        auto p = find(n, ParamType::Int);
        return p ? &p->ints : nullptr;
    }
    const std::vector<float> *findFloats(const std::string &n, ParamType t = ParamType::Float) const {
        auto p = find(n, t);
        return p ? &p->floats : nullptr;
    }
    void reportUnused() const {
        for (auto &q : items)
            if (!q.lookedUp) hostWarning("Parameter \"%s\" not used", q.name.c_str());
    }
    std::vector<Param> items;

  private:
    // the reference keeps one list per storage class; a re-declaration of the
    // same name within a class replaces the earlier one (paramset.cpp:60-108)
    static int cls(ParamType t) {
        switch (t) {
        case ParamType::Int: return 0;
        case ParamType::Bool: return 1;
        case ParamType::Float: return 2;
        case ParamType::Point2: return 3;
        case ParamType::Vector2: return 4;
        case ParamType::Point3: return 5;
        case ParamType::Vector3: return 6;
        case ParamType::Normal: return 7;
        case ParamType::Spectrum: return 8;
        case ParamType::String: return 9;
        default: return 10;
        }
    }
    static bool sameClass(ParamType a, ParamType b) { return cls(a) == cls(b); }
};

}  // namespace pg
#endif
