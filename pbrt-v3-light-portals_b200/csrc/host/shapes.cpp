// Shape factories of the host front end: they turn a `Shape "name" params`
// directive into world-space geometry records (SceneDesc) the way the
// reference's MakeShapes does (src/core/api.cpp:432-546), for the shape types
// on the hot path: trianglemesh, plymesh, loopsubdiv, sphere, aaplane.
#include <cstdio>
#include <cstring>
#include <fstream>
#include <map>
#include <sstream>

#include "api.h"
#include "xform.h"

namespace pg {

// TriangleMesh ctor, shapes/triangle.cpp:55-93: positions, normals and
// tangents go to world space once; uv is copied.
static int addMesh(SceneDesc &sc, const Xform &o2w, bool reverseOrientation, const std::vector<int> &indices,
                   const std::vector<V3> &P, const std::vector<V3> *N, const std::vector<V3> *S,
                   const std::vector<float> *UV) {
    MeshData m;
    m.reverseOrientation = reverseOrientation;
    m.swapsHandedness = xfSwapsHandedness(o2w);
    m.indices = indices;
    m.p.resize(P.size());
    for (size_t i = 0; i < P.size(); ++i) m.p[i] = xformPoint(o2w.m, P[i]);
    if (UV) m.uv = *UV;
    if (N) {
        m.n.resize(N->size());
        for (size_t i = 0; i < N->size(); ++i) m.n[i] = xformNormal(o2w.mInv, (*N)[i]);
    }
    if (S) {
        m.s.resize(S->size());
        for (size_t i = 0; i < S->size(); ++i) m.s[i] = xformVector(o2w.m, (*S)[i]);
    }
    sc.meshes.push_back(std::move(m));
    return (int)sc.meshes.size() - 1;
}

static std::vector<V3> toV3(const std::vector<float> &f) {
    std::vector<V3> r(f.size() / 3);
    for (size_t i = 0; i < r.size(); ++i) r[i] = v3(f[3 * i], f[3 * i + 1], f[3 * i + 2]);
    return r;
}

// CreateTriangleMeshShape, shapes/triangle.cpp:911-1007
static bool makeTriangleMesh(SceneDesc &sc, const Xform &o2w, bool ro, const ParamSet &params, ShapeResult *out) {
    const std::vector<int> *vi = params.findInts("indices");
    const std::vector<float> *P = params.findFloats("P", ParamType::Point3);
    const std::vector<float> *uvs = params.findFloats("uv", ParamType::Point2);
    if (!uvs) uvs = params.findFloats("st", ParamType::Point2);
    if (!uvs) uvs = params.findFloats("uv", ParamType::Float);
    if (!uvs) uvs = params.findFloats("st", ParamType::Float);
    int npi = P ? (int)P->size() / 3 : 0;
    std::vector<float> uvCopy;
    if (uvs) {
        int nuvi = (int)uvs->size() / 2;
        if (nuvi < npi) {
            hostError("Not enough of \"uv\"s for triangle mesh.  Expected %d, found %d.  Discarding.", npi, nuvi);
            uvs = nullptr;
        } else {
            if (nuvi > npi)
                hostWarning("More \"uv\"s provided than will be used for triangle mesh.  (%d expcted, %d found)", npi, nuvi);
            uvCopy.assign(uvs->begin(), uvs->begin() + 2 * npi);
        }
    }
    if (!vi) {
        hostError("Vertex indices \"indices\" not provided with triangle mesh shape");
        return false;
    }
    if (!P) {
        hostError("Vertex positions \"P\" not provided with triangle mesh shape");
        return false;
    }
    const std::vector<float> *S = params.findFloats("S", ParamType::Vector3);
    if (S && (int)S->size() / 3 != npi) {
        hostError("Number of \"S\"s for triangle mesh must match \"P\"s");
        S = nullptr;
    }
    const std::vector<float> *N = params.findFloats("N", ParamType::Normal);
    if (N && (int)N->size() / 3 != npi) {
        hostError("Number of \"N\"s for triangle mesh must match \"P\"s");
        N = nullptr;
    }
    for (int idx : *vi)
        if (idx >= npi) {
            hostError("trianglemesh has out of-bounds vertex index %d (%d \"P\" values were given", idx, npi);
            return false;
        }
    if (params.findTexture("alpha") != "" || params.findTexture("shadowalpha") != "")
        hostWarning("alpha textures are not supported on the GPU path; ignoring");
    params.findOneFloat("alpha", 1.f);
    params.findOneFloat("shadowalpha", 1.f);
    std::vector<V3> Pv = toV3(*P), Nv, Sv;
    if (N) Nv = toV3(*N);
    if (S) Sv = toV3(*S);
    std::vector<int> idx(vi->begin(), vi->begin() + 3 * (vi->size() / 3));
    out->mesh = addMesh(sc, o2w, ro, idx, Pv, N ? &Nv : nullptr, S ? &Sv : nullptr, uvs ? &uvCopy : nullptr);
    out->type = PBRTGPU_PRIM_TRIANGLE;
    out->count = (int)idx.size() / 3;
    return true;
}

// ---- PLY reader (ascii + binary little/big endian); mirrors what
// shapes/plymesh.cpp:176-290 pulls out of rply: x y z [nx ny nz] [u v | s t |
// texture_u texture_v | texture_s texture_t], face vertex_indices (triangles
// and quads; a quad becomes (0,1,2) (3,0,2), plymesh.cpp:132-140).
namespace {
struct PlyProp {
    std::string name, type, countType;
    bool isList = false;
};
struct PlyElem {
    std::string name;
    long count = 0;
    std::vector<PlyProp> props;
};
int plyTypeSize(const std::string &t) {
    if (t == "char" || t == "uchar" || t == "int8" || t == "uint8") return 1;
    if (t == "short" || t == "ushort" || t == "int16" || t == "uint16") return 2;
    if (t == "int" || t == "uint" || t == "float" || t == "int32" || t == "uint32" || t == "float32") return 4;
    if (t == "double" || t == "float64") return 8;
    return 0;
}
double plyDecode(const std::string &t, const unsigned char *b, bool swap) {
    unsigned char tmp[8];
    int n = plyTypeSize(t);
    for (int i = 0; i < n; ++i) tmp[i] = swap ? b[n - 1 - i] : b[i];
    if (t == "char" || t == "int8") return (signed char)tmp[0];
    if (t == "uchar" || t == "uint8") return tmp[0];
    if (t == "short" || t == "int16") { int16_t v; memcpy(&v, tmp, 2); return v; }
    if (t == "ushort" || t == "uint16") { uint16_t v; memcpy(&v, tmp, 2); return v; }
    if (t == "int" || t == "int32") { int32_t v; memcpy(&v, tmp, 4); return v; }
    if (t == "uint" || t == "uint32") { uint32_t v; memcpy(&v, tmp, 4); return v; }
    if (t == "float" || t == "float32") { float v; memcpy(&v, tmp, 4); return v; }
    double v;
    memcpy(&v, tmp, 8);
    return v;
}
}  // namespace

static bool makePLYMesh(SceneDesc &sc, const Xform &o2w, bool ro, const ParamSet &params, ShapeResult *out) {
    std::string filename = resolveFilename(sc, params.findOneString("filename", ""));
    std::ifstream f(filename, std::ios::binary);
    if (!f) {
        hostError("Couldn't open PLY file \"%s\"", filename.c_str());
        return false;
    }
    std::string line;
    std::getline(f, line);
    if (line.substr(0, 3) != "ply") {
        hostError("Unable to read the header of PLY file \"%s\"", filename.c_str());
        return false;
    }
    std::string format;
    std::vector<PlyElem> elems;
    while (std::getline(f, line)) {
        if (!line.empty() && line.back() == '\r') line.pop_back();
        std::istringstream ss(line);
        std::string tok;
        ss >> tok;
        if (tok == "format") ss >> format;
        else if (tok == "element") {
            PlyElem e;
            ss >> e.name >> e.count;
            elems.push_back(e);
        } else if (tok == "property" && !elems.empty()) {
            PlyProp p;
            ss >> p.type;
            if (p.type == "list") {
                p.isList = true;
                ss >> p.countType >> p.type;
            }
            ss >> p.name;
            elems.back().props.push_back(p);
        } else if (tok == "end_header")
            break;
    }
    bool ascii = format == "ascii", swap = format == "binary_big_endian";
    long vertexCount = 0, faceCount = 0;
    for (auto &e : elems) {
        if (e.name == "vertex") vertexCount = e.count;
        if (e.name == "face") faceCount = e.count;
    }
    if (vertexCount == 0 || faceCount == 0) {
        hostError("%s: PLY file is invalid! No face/vertex elements found!", filename.c_str());
        return false;
    }
    std::vector<V3> P(vertexCount), N;
    std::vector<float> UV;
    std::vector<int> indices;
    bool error = false;
    auto readScalar = [&](const std::string &type) -> double {
        if (ascii) {
            double v = 0;
            f >> v;
            return v;
        }
        unsigned char b[8];
        f.read((char *)b, plyTypeSize(type));
        return plyDecode(type, b, swap);
    };
    for (auto &e : elems) {
        // property slots of interest
        int ix = -1, iy = -1, iz = -1, inx = -1, iny = -1, inz = -1, iu = -1, iv = -1;
        if (e.name == "vertex") {
            auto idx = [&](const char *n) {
                for (size_t k = 0; k < e.props.size(); ++k)
                    if (e.props[k].name == n) return (int)k;
                return -1;
            };
            ix = idx("x"), iy = idx("y"), iz = idx("z");
            if (ix < 0 || iy < 0 || iz < 0) {
                hostError("%s: Vertex coordinate property not found!", filename.c_str());
                return false;
            }
            inx = idx("nx"), iny = idx("ny"), inz = idx("nz");
            if (inx >= 0 && iny >= 0 && inz >= 0) N.resize(vertexCount); else inx = -1;
            const char *uvNames[4][2] = {{"u", "v"}, {"s", "t"}, {"texture_u", "texture_v"}, {"texture_s", "texture_t"}};
            for (auto &nm : uvNames)
                if (idx(nm[0]) >= 0 && idx(nm[1]) >= 0) {
                    iu = idx(nm[0]), iv = idx(nm[1]);
                    break;
                }
            if (iu >= 0) UV.resize(2 * vertexCount);
        }
        for (long i = 0; i < e.count; ++i) {
            for (size_t k = 0; k < e.props.size(); ++k) {
                const PlyProp &p = e.props[k];
                if (!p.isList) {
                    float v = (float)readScalar(p.type);
                    if (e.name == "vertex") {
                        if ((int)k == ix) P[i].x = v; else if ((int)k == iy) P[i].y = v; else if ((int)k == iz) P[i].z = v;
                        else if (inx >= 0 && (int)k == inx) N[i].x = v; else if (inx >= 0 && (int)k == iny) N[i].y = v;
                        else if (inx >= 0 && (int)k == inz) N[i].z = v;
                        else if (iu >= 0 && (int)k == iu) UV[2 * i] = v; else if (iu >= 0 && (int)k == iv) UV[2 * i + 1] = v;
                    }
                } else {
                    long len = (long)readScalar(p.countType);
                    int face[4] = {0, 0, 0, 0};
                    for (long j = 0; j < len; ++j) {
                        int value = (int)readScalar(p.type);
                        if (j < 4) face[j] = value;
                        if (e.name == "face" && p.name == "vertex_indices" && (value < 0 || value >= vertexCount) &&
                            (len == 3 || len == 4)) {
                            hostError("plymesh: Vertex reference %i is out of bounds! Valid range is [0..%i)", value, (int)vertexCount);
                            error = true;
                        }
                    }
                    if (e.name == "face" && p.name == "vertex_indices") {
                        if (len != 3 && len != 4) {
                            hostWarning("plymesh: Ignoring face with %i vertices (only triangles and quads are supported!)", (int)len);
                            continue;
                        }
                        indices.push_back(face[0]); indices.push_back(face[1]); indices.push_back(face[2]);
                        if (len == 4) { indices.push_back(face[3]); indices.push_back(face[0]); indices.push_back(face[2]); }
                    }
                }
            }
        }
    }
    if (!f && !ascii) {
        hostError("%s: unable to read the contents of PLY file", filename.c_str());
        return false;
    }
    if (error) return false;
    params.findTexture("alpha"); params.findTexture("shadowalpha");
    params.findOneFloat("alpha", 1.f); params.findOneFloat("shadowalpha", 1.f);
    out->mesh = addMesh(sc, o2w, ro, indices, P, N.empty() ? nullptr : &N, nullptr, UV.empty() ? nullptr : &UV);
    out->type = PBRTGPU_PRIM_TRIANGLE;
    out->count = (int)indices.size() / 3;
    return true;
}

bool makeLoopSubdiv(SceneDesc &sc, const Xform &o2w, bool ro, const ParamSet &params, ShapeResult *out);  // loopsubdiv.cpp

int addMeshForSubdiv(SceneDesc &sc, const Xform &o2w, bool ro, const std::vector<int> &indices, const std::vector<V3> &P,
                     const std::vector<V3> &N) {
    return addMesh(sc, o2w, ro, indices, P, &N, nullptr, nullptr);
}

// CreateSphereShape + Sphere ctor, shapes/sphere.cpp:326-336, sphere.h:50-59
static bool makeSphere(SceneDesc &sc, const Xform &o2w, bool ro, const ParamSet &params, ShapeResult *out) {
    float radius = params.findOneFloat("radius", 1.f);
    float zmin = params.findOneFloat("zmin", -radius);
    float zmax = params.findOneFloat("zmax", radius);
    float phimax = params.findOneFloat("phimax", 360.f);
    SphereData s;
    s.radius = radius;
    s.zMin = clampf(std::min(zmin, zmax), -radius, radius);
    s.zMax = clampf(std::max(zmin, zmax), -radius, radius);
    s.thetaMin = std::acos(clampf(std::min(zmin, zmax) / radius, -1, 1));
    s.thetaMax = std::acos(clampf(std::max(zmin, zmax) / radius, -1, 1));
    s.phiMax = radians(clampf(phimax, 0, 360));
    s.reverseOrientation = ro;
    s.swapsHandedness = xfSwapsHandedness(o2w);
    s.o2w = o2w;
    sc.spheres.push_back(s);
    out->type = PBRTGPU_PRIM_SPHERE;
    out->mesh = (int)sc.spheres.size() - 1;
    out->count = 1;
    return true;
}

// CreateAAPlaneShape + AAPlaneShape ctor, shapes/plane.cpp:117-128, plane.h:15-32.
// The "facingFw" parameter is read and then overwritten by !reverseOrientation.
static bool makeAAPlane(SceneDesc &sc, const Xform &o2w, bool ro, const ParamSet &params, ShapeResult *out) {
    PlaneData p;
    p.lo = params.findOnePoint3("lo", v3(0, 0, 0));
    p.hi = params.findOnePoint3("hi", v3(0, 0, 0));
    p.ax = params.findOneInt("axis", 2);
    params.findOneBool("facingFw", true);
    p.facingFw = !ro;
    p.reverseOrientation = ro;
    p.swapsHandedness = xfSwapsHandedness(o2w);
    p.o2w = o2w;
    int ax0 = p.ax == 2 ? 0 : (p.ax == 0 ? 1 : 2), ax1 = p.ax == 2 ? 1 : (p.ax == 0 ? 2 : 0);
    V3 loW = xformPoint(o2w.m, p.lo), hiW = xformPoint(o2w.m, p.hi);
    p.area = (hiW[ax0] - loW[ax0]) * (hiW[ax1] - loW[ax1]);
    sc.planes.push_back(p);
    out->type = PBRTGPU_PRIM_AAPLANE;
    out->mesh = (int)sc.planes.size() - 1;
    out->count = 1;
    return true;
}

bool makeShape(SceneDesc &sc, const std::string &name, const Xform &o2w, bool reverseOrientation,
               const ParamSet &params, ShapeResult *out) {
    if (name == "trianglemesh") return makeTriangleMesh(sc, o2w, reverseOrientation, params, out);
    if (name == "plymesh") return makePLYMesh(sc, o2w, reverseOrientation, params, out);
    if (name == "loopsubdiv") return makeLoopSubdiv(sc, o2w, reverseOrientation, params, out);
    if (name == "sphere") return makeSphere(sc, o2w, reverseOrientation, params, out);
    if (name == "aaplane") return makeAAPlane(sc, o2w, reverseOrientation, params, out);
    hostError("Shape \"%s\" is not supported by the GPU path integrator (supported: trianglemesh, plymesh, "
              "loopsubdiv, sphere, aaplane)", name.c_str());
    return false;
}

// Shape::WorldBound per primitive kind: Triangle::WorldBound (triangle.cpp:179-187),
// Shape::WorldBound = ObjectToWorld(ObjectBound()) for sphere / aaplane (shape.cpp:54).
Bounds3 primWorldBound(const SceneDesc &s, const PrimDesc &p) {
    if (p.type == PBRTGPU_PRIM_TRIANGLE) {
        const MeshData &m = s.meshes[p.shape];
        V3 p0 = m.p[m.indices[3 * p.tri]], p1 = m.p[m.indices[3 * p.tri + 1]], p2 = m.p[m.indices[3 * p.tri + 2]];
        return unionB(boundsOf(p0, p1), p2);
    }
    if (p.type == PBRTGPU_PRIM_SPHERE) {
        const SphereData &sp = s.spheres[p.shape];
        Bounds3 ob = boundsOf(v3(-sp.radius, -sp.radius, sp.zMin), v3(sp.radius, sp.radius, sp.zMax));
        return xfBounds(sp.o2w, ob);
    }
    const PlaneData &pl = s.planes[p.shape];
    Bounds3 ob;  // AAPlaneShape::ObjectBound returns {lo, hi} (plane.cpp:6-8) -> Bounds3(p1,p2) takes min/max
    ob = boundsOf(pl.lo, pl.hi);
    return xfBounds(pl.o2w, ob);
}

}  // namespace pg
