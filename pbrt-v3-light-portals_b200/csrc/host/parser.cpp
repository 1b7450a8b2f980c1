// .pbrt scene-file reader: tokenizer + directive dispatch into SceneApi.
// Grammar, parameter typing and number parsing follow src/core/parser.cpp of
// the reference (tokens: quoted strings with escapes, '[' ']', '#' comments,
// bare words; numbers through strtol/strtof so "1.5," parses as 1.5;
// "type name" declarations; single values without brackets; Include).
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <functional>
#include <memory>
#include <sstream>

#include "api.h"

namespace pg {

namespace {

struct Tokenizer {
    std::string text, filename;
    size_t pos = 0;
    int line = 1;
    bool error = false;

    // returns false at EOF; tok keeps quotes for quoted strings (parser.cpp:262-320)
    bool next(std::string &tok) {
        while (pos < text.size()) {
            char ch = text[pos++];
            if (ch == '\n') { ++line; continue; }
            if (ch == ' ' || ch == '\t' || ch == '\r') continue;
            if (ch == '"') {
                tok = "\"";
                while (true) {
                    if (pos >= text.size()) { fail("premature EOF"); return false; }
                    ch = text[pos++];
                    if (ch == '"') break;
                    if (ch == '\n') { fail("unterminated string"); return false; }
                    if (ch == '\\') {
                        if (pos >= text.size()) { fail("premature EOF"); return false; }
                        char e = text[pos++];
                        switch (e) {
                        case 'b': ch = '\b'; break;
                        case 'f': ch = '\f'; break;
                        case 'n': ch = '\n'; break;
                        case 'r': ch = '\r'; break;
                        case 't': ch = '\t'; break;
                        case '\\': ch = '\\'; break;
                        case '\'': ch = '\''; break;
                        case '"': ch = '"'; break;
                        default: fail("unexpected escaped character"); return false;
                        }
                    }
                    tok.push_back(ch);
                }
                tok.push_back('"');
                return true;
            }
            if (ch == '[' || ch == ']') { tok = std::string(1, ch); return true; }
            if (ch == '#') {
                while (pos < text.size() && text[pos] != '\n' && text[pos] != '\r') ++pos;
                continue;
            }
            tok = std::string(1, ch);
            while (pos < text.size()) {
                ch = text[pos];
                if (ch == ' ' || ch == '\n' || ch == '\t' || ch == '\r' || ch == '"' || ch == '[' || ch == ']') break;
                tok.push_back(ch);
                ++pos;
            }
            return true;
        }
        return false;
    }
    void fail(const char *msg) {
        hostError("%s:%d: %s", filename.c_str(), line, msg);
        error = true;
    }
};

bool isQuoted(const std::string &s) { return s.size() >= 2 && s.front() == '"' && s.back() == '"'; }
std::string dequote(const std::string &s) { return s.substr(1, s.size() - 2); }

struct ParseError {};

// parseNumber, parser.cpp:322-368
double parseNumber(const std::string &str) {
    if (str.size() == 1) {
        if (!(str[0] >= '0' && str[0] <= '9')) {
            hostError("\"%c\": expected a number", str[0]);
            throw ParseError();
        }
        return str[0] - '0';
    }
    bool isInteger = true;
    for (char ch : str)
        if (!(ch >= '0' && ch <= '9')) isInteger = false;
    char *endptr = nullptr;
    double val;
    if (isInteger) val = double(strtol(str.c_str(), &endptr, 10));
    else val = strtof(str.c_str(), &endptr);
    if (val == 0 && endptr == str.c_str()) {
        hostError("%s: expected a number", str.c_str());
        throw ParseError();
    }
    return val;
}

struct Parser {
    SceneApi &api;
    SceneDesc &sc;
    std::vector<std::unique_ptr<Tokenizer>> stack;
    bool haveUnget = false;
    std::string ungetTok;

    Parser(SceneApi &a, SceneDesc &s) : api(a), sc(s) {}

    bool nextToken(std::string &tok, bool required) {
        if (haveUnget) {
            haveUnget = false;
            tok = ungetTok;
            return true;
        }
        while (!stack.empty()) {
            if (stack.back()->next(tok)) return true;
            if (stack.back()->error) throw ParseError();
            stack.pop_back();
        }
        if (required) {
            hostError("premature EOF");
            throw ParseError();
        }
        return false;
    }
    void unget(const std::string &t) {
        haveUnget = true;
        ungetTok = t;
    }
    std::string quotedString() {
        std::string t;
        nextToken(t, true);
        if (!isQuoted(t)) {
            hostError("\"%s\": expected quoted string", t.c_str());
            throw ParseError();
        }
        return dequote(t);
    }
    void numbers(float *v, int n, bool brackets) {
        std::string t;
        if (brackets) {
            nextToken(t, true);
            if (t != "[") { hostError("syntax error: expected \"[\" but got \"%s\"", t.c_str()); throw ParseError(); }
        }
        for (int i = 0; i < n; ++i) {
            nextToken(t, true);
            v[i] = (float)parseNumber(t);
        }
        if (brackets) {
            nextToken(t, true);
            if (t != "]") { hostError("syntax error: expected \"]\" but got \"%s\"", t.c_str()); throw ParseError(); }
        }
    }

    bool lookupType(const std::string &decl, ParamType *type, std::string *kind, std::string *name) {
        std::istringstream ss(decl);
        std::string t, n;
        ss >> t >> n;
        if (t.empty()) { hostError("Parameter \"%s\" doesn't have a type declaration?!", decl.c_str()); return false; }
        if (n.empty()) { hostError("Unable to find parameter name from \"%s\"", decl.c_str()); return false; }
        *kind = t;
        *name = n;
        if (t == "float") *type = ParamType::Float;
        else if (t == "integer") *type = ParamType::Int;
        else if (t == "bool") *type = ParamType::Bool;
        else if (t == "point2") *type = ParamType::Point2;
        else if (t == "vector2") *type = ParamType::Vector2;
        else if (t == "point3" || t == "point") *type = ParamType::Point3;
        else if (t == "vector3" || t == "vector") *type = ParamType::Vector3;
        else if (t == "normal") *type = ParamType::Normal;
        else if (t == "string") *type = ParamType::String;
        else if (t == "texture") *type = ParamType::Texture;
        else if (t == "color" || t == "rgb" || t == "xyz" || t == "blackbody" || t == "spectrum") *type = ParamType::Spectrum;
        else { hostError("Unable to decode type from \"%s\"", decl.c_str()); return false; }
        return true;
    }

    // parseParams + AddParam, parser.cpp:470-760
    ParamSet params() {
        ParamSet ps;
        std::string decl;
        while (nextToken(decl, false)) {
            if (!isQuoted(decl)) {
                unget(decl);
                break;
            }
            std::vector<double> nums;
            std::vector<std::string> strs;
            auto addVal = [&](const std::string &val) {
                if (isQuoted(val)) {
                    if (!nums.empty()) { hostError("mixed string and numeric parameters"); throw ParseError(); }
                    strs.push_back(dequote(val));
                } else {
                    if (!strs.empty()) { hostError("mixed string and numeric parameters"); throw ParseError(); }
                    nums.push_back(parseNumber(val));
                }
            };
            std::string val;
            nextToken(val, true);
            if (val == "[") {
                while (true) {
                    nextToken(val, true);
                    if (val == "]") break;
                    addVal(val);
                }
            } else
                addVal(val);

            Param p;
            std::string kind;
            if (!lookupType(dequote(decl), &p.type, &kind, &p.name)) {
                hostWarning("Type of parameter \"%s\" is unknown", dequote(decl).c_str());
                continue;
            }
            bool wantsString = p.type == ParamType::String || p.type == ParamType::Texture || p.type == ParamType::Bool ||
                               (kind == "spectrum" && !strs.empty());
            if (wantsString && strs.empty()) { hostError("Expected string parameter value for parameter \"%s\" with type \"%s\". Ignoring.", p.name.c_str(), kind.c_str()); continue; }
            if (!wantsString && nums.empty() && !strs.empty()) { hostError("Expected numeric parameter value for parameter \"%s\" with type \"%s\". Ignoring.", p.name.c_str(), kind.c_str()); continue; }
            size_t n = nums.size();
            switch (p.type) {
            case ParamType::Int:
                for (double d : nums) p.ints.push_back((int)d);
                break;
            case ParamType::Bool:
                for (auto &s : strs) {
                    if (s == "true") p.bools.push_back(1);
                    else if (s == "false") p.bools.push_back(0);
                    else {
                        hostWarning("Value \"%s\" unknown for Boolean parameter \"%s\".Using \"false\".", s.c_str(), p.name.c_str());
                        p.bools.push_back(0);
                    }
                }
                break;
            case ParamType::Float:
                for (double d : nums) p.floats.push_back((float)d);
                break;
            case ParamType::Point2:
            case ParamType::Vector2:
                if (n % 2) hostWarning("Excess values given with %s parameter \"%s\". Ignoring last one of them.", kind.c_str(), p.name.c_str());
                for (size_t i = 0; i < n - n % 2; ++i) p.floats.push_back((float)nums[i]);
                break;
            case ParamType::Point3:
            case ParamType::Vector3:
            case ParamType::Normal:
                if (n % 3) hostWarning("Excess values given with %s parameter \"%s\". Ignoring last %d of them.", kind.c_str(), p.name.c_str(), int(n % 3));
                for (size_t i = 0; i < n - n % 3; ++i) p.floats.push_back((float)nums[i]);
                break;
            case ParamType::Spectrum:
                if (kind == "color" || kind == "rgb") {
                    if (n % 3) hostWarning("Excess RGB values given with parameter \"%s\". Ignoring last %d of them", p.name.c_str(), int(n % 3));
                    for (size_t i = 0; i + 2 < n; i += 3) p.spectra.push_back(RGB((float)nums[i], (float)nums[i + 1], (float)nums[i + 2]));
                } else if (kind == "xyz") {
                    if (n % 3) hostWarning("Excess XYZ values given with parameter \"%s\". Ignoring last %d of them", p.name.c_str(), int(n % 3));
                    for (size_t i = 0; i + 2 < n; i += 3) {
                        float xyz[3] = {(float)nums[i], (float)nums[i + 1], (float)nums[i + 2]};
                        p.spectra.push_back(rgbFromXYZ(xyz));
                    }
                } else if (kind == "blackbody") {
                    if (n % 2) hostWarning("Excess value given with blackbody parameter \"%s\". Ignoring extra one.", p.name.c_str());
                    for (size_t i = 0; i + 1 < n; i += 2) p.spectra.push_back(rgbFromBlackbody((float)nums[i], (float)nums[i + 1]));
                } else if (!strs.empty()) {
                    for (auto &fn : strs) p.spectra.push_back(spectrumFromFile(fn));
                } else {
                    if (n % 2) hostWarning("Non-even number of values given with sampled spectrum parameter \"%s\". Ignoring extra.", p.name.c_str());
                    std::vector<float> wl, v;
                    for (size_t i = 0; i + 1 < n; i += 2) {
                        wl.push_back((float)nums[i]);
                        v.push_back((float)nums[i + 1]);
                    }
                    p.spectra.push_back(rgbFromSampled(wl.data(), v.data(), (int)wl.size()));
                }
                break;
            case ParamType::String:
                p.strings = strs;
                break;
            case ParamType::Texture:
                if (strs.size() == 1) p.strings = strs;
                else { hostError("Only one string allowed for \"texture\" parameter \"%s\"", p.name.c_str()); continue; }
                break;
            }
            ps.add(std::move(p));
        }
        return ps;
    }

    // ParamSet::AddSampledSpectrumFiles, paramset.cpp:167-211: "lambda value" pairs
    RGB spectrumFromFile(const std::string &fn) {
        std::string path = resolveFilename(sc, fn);
        std::ifstream f(path);
        if (!f) {
            hostWarning("Unable to read SPD file \"%s\".  Using black distribution.", path.c_str());
            return RGB(0.f);
        }
        std::vector<float> vals;
        std::string line;
        while (std::getline(f, line)) {
            size_t h = line.find('#');
            if (h != std::string::npos) line.resize(h);
            std::istringstream ss(line);
            float v;
            while (ss >> v) vals.push_back(v);
        }
        if (vals.size() % 2) {
            hostWarning("Extra value found in spectrum file \"%s\". Ignoring it.", path.c_str());
            vals.pop_back();
        }
        std::vector<float> wl, v;
        for (size_t i = 0; i + 1 < vals.size(); i += 2) wl.push_back(vals[i]), v.push_back(vals[i + 1]);
        return rgbFromSampled(wl.data(), v.data(), (int)wl.size());
    }

    bool pushFile(const std::string &path) {
        std::ifstream f(path, std::ios::binary);
        if (!f) {
            hostError("%s: unable to open file", path.c_str());
            return false;
        }
        std::stringstream ss;
        ss << f.rdbuf();
        auto t = std::make_unique<Tokenizer>();
        t->text = ss.str();
        t->filename = path;
        stack.push_back(std::move(t));
        return true;
    }

    void run() {
        std::string tok;
        while (nextToken(tok, false)) {
            auto named = [&](void (SceneApi::*fn)(const std::string &, const ParamSet &)) {
                std::string n = quotedString();
                ParamSet ps = params();
                (api.*fn)(n, ps);
            };
            float v[16];
            if (tok == "AttributeBegin") api.AttributeBegin();
            else if (tok == "AttributeEnd") api.AttributeEnd();
            else if (tok == "ActiveTransform") { std::string a; nextToken(a, true); api.ActiveTransform(a); }
            else if (tok == "AreaLightSource") named(&SceneApi::AreaLightSource);
            else if (tok == "Accelerator") named(&SceneApi::Accelerator);
            else if (tok == "ConcatTransform") { numbers(v, 16, true); api.ConcatTransform(v); }
            else if (tok == "CoordinateSystem") api.CoordinateSystem(quotedString());
            else if (tok == "CoordSysTransform") api.CoordSysTransform(quotedString());
            else if (tok == "Camera") named(&SceneApi::Camera);
            else if (tok == "Film") named(&SceneApi::Film);
            else if (tok == "Integrator") named(&SceneApi::Integrator);
            else if (tok == "Include") {
                std::string fn = resolveFilename(sc, quotedString());
                if (!pushFile(fn)) throw ParseError();
            } else if (tok == "Identity") api.Identity();
            else if (tok == "LightSource") named(&SceneApi::LightSource);
            else if (tok == "LookAt") { numbers(v, 9, false); api.LookAt(v[0], v[1], v[2], v[3], v[4], v[5], v[6], v[7], v[8]); }
            else if (tok == "MakeNamedMaterial") named(&SceneApi::MakeNamedMaterial);
            else if (tok == "MakeNamedMedium") named(&SceneApi::MakeNamedMedium);
            else if (tok == "Material") named(&SceneApi::Material);
            else if (tok == "MediumInterface") {
                std::string a = quotedString(), b = a, t;
                if (nextToken(t, false)) { if (isQuoted(t)) b = dequote(t); else unget(t); }
                api.MediumInterface(a, b);
            } else if (tok == "NamedMaterial") api.NamedMaterial(quotedString());
            else if (tok == "ObjectBegin") api.ObjectBegin(quotedString());
            else if (tok == "ObjectEnd") api.ObjectEnd();
            else if (tok == "ObjectInstance") api.ObjectInstance(quotedString());
            else if (tok == "PixelFilter") named(&SceneApi::PixelFilter);
            else if (tok == "ReverseOrientation") api.ReverseOrientation();
            else if (tok == "Rotate") { numbers(v, 4, false); api.Rotate(v[0], v[1], v[2], v[3]); }
            else if (tok == "Shape") named(&SceneApi::Shape);
            else if (tok == "Sampler") named(&SceneApi::Sampler);
            else if (tok == "Scale") { numbers(v, 3, false); api.Scale(v[0], v[1], v[2]); }
            else if (tok == "TransformBegin") api.TransformBegin();
            else if (tok == "TransformEnd") api.TransformEnd();
            else if (tok == "Transform") { numbers(v, 16, true); api.Transform(v); }
            else if (tok == "Translate") { numbers(v, 3, false); api.Translate(v[0], v[1], v[2]); }
            else if (tok == "TransformTimes") { numbers(v, 2, false); api.TransformTimes(v[0], v[1]); }
            else if (tok == "Texture") {
                std::string name = quotedString(), type = quotedString(), texName = quotedString();
                ParamSet ps = params();
                api.Texture(name, type, texName, ps);
            } else if (tok == "WorldBegin") api.WorldBegin();
            else if (tok == "WorldEnd") api.WorldEnd();
            else {
                hostError("%s: unknown directive", tok.c_str());
                throw ParseError();
            }
        }
    }
};

}  // namespace

static std::string dirOf(const std::string &path) {
    size_t s = path.find_last_of('/');
    return s == std::string::npos ? std::string(".") : path.substr(0, s);
}

bool parseFile(const std::string &filename, SceneApi &api, SceneDesc &sc) {
    sc.sceneDir = dirOf(filename);
    Parser p(api, sc);
    if (!p.pushFile(filename)) return false;
    try {
        p.run();
    } catch (ParseError &) {
        return false;
    }
    return api.ok();
}

bool parseString(const std::string &text, SceneApi &api, SceneDesc &sc) {
    Parser p(api, sc);
    auto t = std::make_unique<Tokenizer>();
    t->text = text;
    t->filename = "<string>";
    p.stack.push_back(std::move(t));
    try {
        p.run();
    } catch (ParseError &) {
        return false;
    }
    return api.ok();
}

}  // namespace pg
