// nori/object.h — NoriObject, the object factory and the property list.
// Reference: include/nori/object.h:31-161 (class types, lifecycle, NORI_REGISTER_CLASS),
// src/object.cpp:25-38 (lazily created registry), include/nori/proplist.h:31-137,
// src/proplist.cpp:23-60 (typed accessors and their two error messages).
#pragma once
#include "transform.h"
#include <functional>
#include <map>

NORI_NAMESPACE_BEGIN

class PropertyList {
public:
    void setBoolean(const std::string &name, bool v);
    void setInteger(const std::string &name, int v);
    void setFloat(const std::string &name, float v);
    void setString(const std::string &name, const std::string &v);
    void setColor(const std::string &name, const Color3f &v);
    void setPoint(const std::string &name, const Point3f &v);
    void setVector(const std::string &name, const Vector3f &v);
    void setTransform(const std::string &name, const Transform &v);

    bool getBoolean(const std::string &name) const;
    bool getBoolean(const std::string &name, bool def) const;
    int getInteger(const std::string &name) const;
    int getInteger(const std::string &name, int def) const;
    float getFloat(const std::string &name) const;
    float getFloat(const std::string &name, float def) const;
    std::string getString(const std::string &name) const;
    std::string getString(const std::string &name, const std::string &def) const;
    Color3f getColor(const std::string &name) const;
    Color3f getColor(const std::string &name, const Color3f &def) const;
    Point3f getPoint(const std::string &name) const;
    Point3f getPoint(const std::string &name, const Point3f &def) const;
    Vector3f getVector(const std::string &name) const;
    Vector3f getVector(const std::string &name, const Vector3f &def) const;
    Transform getTransform(const std::string &name) const;
    Transform getTransform(const std::string &name, const Transform &def) const;

private:
    enum Type { TBoolean, TInteger, TFloat, TString, TColor, TPoint, TVector, TTransform };
    struct Property {
        Type type = TBoolean;
        bool b = false; int i = 0; float f = 0; std::string s; Vector3f v; Transform t;   // color / point / vector share v
    };
    /// nullptr: not present (caller falls back to the default); throws on a type mismatch
    const Property *find(const std::string &name, Type type, const char *xmlName, bool required) const;
    Property &put(const std::string &name, Type type);
    std::map<std::string, Property> m_properties;
};

/// Base class of everything that can be named in a scene file.
class NoriObject {
public:
    /// Same values as the reference (they double as XML tag ids in the parser).
    enum EClassType {
        EScene = 0, EMesh, EBSDF, EPhaseFunction, EEmitter, EMedium, ECamera, EIntegrator, ESampler, ETest,
        EReconstructionFilter, EClassTypeCount
    };
    virtual ~NoriObject() {}
    virtual EClassType getClassType() const = 0;
    /// Default: children are not accepted (include/nori/object.h:72, src/object.cpp:25-29)
    virtual void addChild(NoriObject *child);
    virtual void setParent(NoriObject *) {}
    /// Called once all properties and children are in place
    virtual void activate() {}
    virtual std::string toString() const = 0;

    static std::string classTypeName(EClassType type) {
        switch (type) {
            case EScene: return "scene"; case EMesh: return "mesh"; case EBSDF: return "bsdf";
            case EEmitter: return "emitter"; case ECamera: return "camera"; case EIntegrator: return "integrator";
            case ESampler: return "sampler"; case ETest: return "test";
            case EPhaseFunction: return "phase"; case EMedium: return "medium";
            case EReconstructionFilter: return "rfilter";
            default: return "<unknown>";
        }
    }
};

class NoriObjectFactory {
public:
    typedef std::function<NoriObject *(const PropertyList &)> Constructor;
    /// Called by NORI_REGISTER_CLASS at static-initialisation time
    static void registerClass(const std::string &name, const Constructor &constr);
    /// Throws NoriException for an unknown name (include/nori/object.h:142-147)
    static NoriObject *createInstance(const std::string &name, const PropertyList &propList);
    static bool isRegistered(const std::string &name);
private:
    static std::map<std::string, Constructor> &registry();
};

/// Registers `cls` under the scene-file name `name` (include/nori/object.h:153-161).
#define NORI_REGISTER_CLASS(cls, name) \
    static nori::NoriObject *cls##_create(const nori::PropertyList &list) { return new cls(list); } \
    static struct cls##_registrar { \
        cls##_registrar() { nori::NoriObjectFactory::registerClass(name, cls##_create); } \
    } cls##_registrar_instance;

/// Parse a scene description and instantiate it (include/nori/parser.h:29; src/parser.cpp).
NoriObject *loadFromXML(const std::string &filename);

NORI_NAMESPACE_END
