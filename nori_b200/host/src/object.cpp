// object.cpp — object factory and property list (see nori/object.h).
#include <nori/object.h>

NORI_NAMESPACE_BEGIN

void NoriObject::addChild(NoriObject *) {
    throw NoriException("NoriObject::addChild() is not implemented for objects of type '%s'!", classTypeName(getClassType()));
}

std::map<std::string, NoriObjectFactory::Constructor> &NoriObjectFactory::registry() {
    // heap-allocated on first use: registration happens from static initialisers of other
    // translation units, in unspecified order
    static std::map<std::string, Constructor> *r = new std::map<std::string, Constructor>();
    return *r;
}

void NoriObjectFactory::registerClass(const std::string &name, const Constructor &constr) { registry()[name] = constr; }

bool NoriObjectFactory::isRegistered(const std::string &name) { return registry().count(name) != 0; }

NoriObject *NoriObjectFactory::createInstance(const std::string &name, const PropertyList &propList) {
    auto it = registry().find(name);
    if (it == registry().end())
        throw NoriException("A constructor for class \"%s\" could not be found!", name);
    return it->second(propList);
}

// ---- PropertyList -----------------------------------------------------------------------------
PropertyList::Property &PropertyList::put(const std::string &name, Type type) {
    auto it = m_properties.find(name);
    if (it != m_properties.end())
        fprintf(stderr, "Property \"%s\" was specified multiple times!\n", name.c_str());
    Property &p = m_properties[name];
    p = Property();
    p.type = type;
    return p;
}

const PropertyList::Property *PropertyList::find(const std::string &name, Type type, const char *xmlName, bool required) const {
    auto it = m_properties.find(name);
    if (it == m_properties.end()) {
        if (required) throw NoriException("Property '%s' is missing!", name);
        return nullptr;
    }
    if (it->second.type != type)
        throw NoriException("Property '%s' has the wrong type! (expected <%s>)!", name, xmlName);
    return &it->second;
}

void PropertyList::setBoolean(const std::string &n, bool v) { put(n, TBoolean).b = v; }
void PropertyList::setInteger(const std::string &n, int v) { put(n, TInteger).i = v; }
void PropertyList::setFloat(const std::string &n, float v) { put(n, TFloat).f = v; }
void PropertyList::setString(const std::string &n, const std::string &v) { put(n, TString).s = v; }
void PropertyList::setColor(const std::string &n, const Color3f &v) { put(n, TColor).v = Vector3f(v.r, v.g, v.b); }
void PropertyList::setPoint(const std::string &n, const Point3f &v) { put(n, TPoint).v = v; }
void PropertyList::setVector(const std::string &n, const Vector3f &v) { put(n, TVector).v = v; }
void PropertyList::setTransform(const std::string &n, const Transform &v) { put(n, TTransform).t = v; }

bool PropertyList::getBoolean(const std::string &n) const { return find(n, TBoolean, "boolean", true)->b; }
bool PropertyList::getBoolean(const std::string &n, bool d) const { auto p = find(n, TBoolean, "boolean", false); return p ? p->b : d; }
int PropertyList::getInteger(const std::string &n) const { return find(n, TInteger, "integer", true)->i; }
int PropertyList::getInteger(const std::string &n, int d) const { auto p = find(n, TInteger, "integer", false); return p ? p->i : d; }
float PropertyList::getFloat(const std::string &n) const { return find(n, TFloat, "float", true)->f; }
float PropertyList::getFloat(const std::string &n, float d) const { auto p = find(n, TFloat, "float", false); return p ? p->f : d; }
std::string PropertyList::getString(const std::string &n) const { return find(n, TString, "string", true)->s; }
std::string PropertyList::getString(const std::string &n, const std::string &d) const { auto p = find(n, TString, "string", false); return p ? p->s : d; }
Color3f PropertyList::getColor(const std::string &n) const { auto p = find(n, TColor, "color", true); return Color3f(p->v.x, p->v.y, p->v.z); }
Color3f PropertyList::getColor(const std::string &n, const Color3f &d) const { auto p = find(n, TColor, "color", false); return p ? Color3f(p->v.x, p->v.y, p->v.z) : d; }
Point3f PropertyList::getPoint(const std::string &n) const { return find(n, TPoint, "point", true)->v; }
Point3f PropertyList::getPoint(const std::string &n, const Point3f &d) const { auto p = find(n, TPoint, "point", false); return p ? p->v : d; }
Vector3f PropertyList::getVector(const std::string &n) const { return find(n, TVector, "vector", true)->v; }
Vector3f PropertyList::getVector(const std::string &n, const Vector3f &d) const { auto p = find(n, TVector, "vector", false); return p ? p->v : d; }
Transform PropertyList::getTransform(const std::string &n) const { return find(n, TTransform, "transform", true)->t; }
Transform PropertyList::getTransform(const std::string &n, const Transform &d) const { auto p = find(n, TTransform, "transform", false); return p ? p->t : d; }

NORI_NAMESPACE_END
