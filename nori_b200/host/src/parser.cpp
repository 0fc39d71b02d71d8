// parser.cpp — scene description -> object graph.
// Reference: src/parser.cpp:29-317. Same tags, same attribute checks, same construction order
// (children are built and activated before their parent; the object is created from the factory by
// its "type" attribute, must report the class type its tag names, then receives addChild/setParent
// for every child object and finally activate()).
#include <nori/object.h>
#include "xml.h"
#include <set>

NORI_NAMESPACE_BEGIN

namespace {

// Tag ids: object classes share NoriObject::EClassType, properties follow.
enum Tag {
    TBoolean = NoriObject::EClassTypeCount, TInteger, TFloat, TString, TPoint, TVector, TColor,
    TTransform, TTranslate, TMatrix, TRotate, TScale, TLookAt, TInvalid
};

const std::map<std::string, int> &tagTable() {
    static const std::map<std::string, int> t = {
        {"scene", NoriObject::EScene}, {"mesh", NoriObject::EMesh}, {"bsdf", NoriObject::EBSDF},
        {"emitter", NoriObject::EEmitter}, {"camera", NoriObject::ECamera}, {"medium", NoriObject::EMedium},
        {"phase", NoriObject::EPhaseFunction}, {"integrator", NoriObject::EIntegrator}, {"sampler", NoriObject::ESampler},
        {"rfilter", NoriObject::EReconstructionFilter}, {"test", NoriObject::ETest},
        {"boolean", TBoolean}, {"integer", TInteger}, {"float", TFloat}, {"string", TString}, {"point", TPoint},
        {"vector", TVector}, {"color", TColor}, {"transform", TTransform}, {"translate", TTranslate},
        {"matrix", TMatrix}, {"rotate", TRotate}, {"scale", TScale}, {"lookat", TLookAt}};
    return t;
}

struct SceneBuilder {
    const std::string &filename;
    const XmlDocument &doc;
    TransformBuilder transform;                      // contents of the <transform> being read

    [[noreturn]] void fail(const XmlNode &n, const std::string &what) const {
        throw NoriException("Error while parsing \"%s\": %s (at %s)", filename, what, doc.where(n.offset));
    }

    /// Every attribute must be expected and every expected attribute present.
    void checkAttributes(const XmlNode &n, std::set<std::string> expected) const {
        for (const auto &a : n.attributes) {
            auto it = expected.find(a.first);
            if (it == expected.end()) fail(n, "unexpected attribute \"" + a.first + "\" in \"" + n.name + "\"");
            expected.erase(it);
        }
        if (!expected.empty()) fail(n, "missing attribute \"" + *expected.begin() + "\" in \"" + n.name + "\"");
    }

    const std::string &attr(const XmlNode &n, const char *key) const { return *n.attribute(key); }

    void setProperty(const XmlNode &n, int tag, PropertyList &list) {
        switch (tag) {
            case TString: checkAttributes(n, {"name", "value"}); list.setString(attr(n, "name"), attr(n, "value")); break;
            case TFloat: checkAttributes(n, {"name", "value"}); list.setFloat(attr(n, "name"), toFloat(attr(n, "value"))); break;
            case TInteger: checkAttributes(n, {"name", "value"}); list.setInteger(attr(n, "name"), toInt(attr(n, "value"))); break;
            case TBoolean: checkAttributes(n, {"name", "value"}); list.setBoolean(attr(n, "name"), toBool(attr(n, "value"))); break;
            case TPoint: checkAttributes(n, {"name", "value"}); list.setPoint(attr(n, "name"), toVector3f(attr(n, "value"))); break;
            case TVector: checkAttributes(n, {"name", "value"}); list.setVector(attr(n, "name"), toVector3f(attr(n, "value"))); break;
            case TColor: {
                checkAttributes(n, {"name", "value"});
                const Vector3f v = toVector3f(attr(n, "value"));
                list.setColor(attr(n, "name"), Color3f(v.x, v.y, v.z));
                break;
            }
            case TTransform: checkAttributes(n, {"name"}); list.setTransform(attr(n, "name"), Transform(transform.result())); break;
            case TTranslate: checkAttributes(n, {"value"}); transform.translate(toVector3f(attr(n, "value"))); break;
            case TScale: checkAttributes(n, {"value"}); transform.scale(toVector3f(attr(n, "value"))); break;
            case TRotate: checkAttributes(n, {"angle", "axis"}); transform.rotate(toFloat(attr(n, "angle")), toVector3f(attr(n, "axis"))); break;
            case TMatrix: {
                checkAttributes(n, {"value"});
                const std::vector<std::string> tok = tokenize(attr(n, "value"));
                if (tok.size() != 16) throw NoriException("Expected 16 values");
                float v[16];
                for (int i = 0; i < 16; ++i) v[i] = toFloat(tok[(size_t) i]);
                transform.matrix(v);
                break;
            }
            case TLookAt:
                checkAttributes(n, {"origin", "target", "up"});
                transform.lookAt(toVector3f(attr(n, "origin")), toVector3f(attr(n, "target")), toVector3f(attr(n, "up")));
                break;
            default: throw NoriException("Unhandled element \"%s\"", n.name);
        }
    }

    NoriObject *build(const XmlNode &n, PropertyList &parentList, int parentTag) {
        if (n.kind == XmlNode::Comment || n.kind == XmlNode::Declaration) return nullptr;
        if (n.kind != XmlNode::Element) fail(n, "unexpected content");
        auto it = tagTable().find(n.name);
        if (it == tagTable().end()) fail(n, "unexpected tag \"" + n.name + "\"");
        const int tag = it->second;

        const bool hasParent = parentTag != TInvalid;
        const bool parentIsObject = hasParent && parentTag < NoriObject::EClassTypeCount;
        const bool isObject = tag < NoriObject::EClassTypeCount;
        const bool parentIsTransform = parentTag == TTransform;
        const bool isTransformOp = tag == TTranslate || tag == TRotate || tag == TScale || tag == TLookAt || tag == TMatrix;
        if (!hasParent && !isObject) fail(n, "root element \"" + n.name + "\" must be a Nori object");
        if (parentIsTransform != isTransformOp) fail(n, "transform nodes can only contain transform operations");
        if (hasParent && !parentIsObject && !(parentIsTransform && isTransformOp))
            fail(n, "node \"" + n.name + "\" requires a Nori object as parent");

        if (tag == TTransform) transform = TransformBuilder();

        PropertyList props;
        std::vector<NoriObject *> children;
        for (const XmlNode &ch : n.children)
            if (NoriObject *child = build(ch, props, tag)) children.push_back(child);

        NoriObject *result = nullptr;
        try {
            if (isObject) {
                std::string type;
                if (tag == NoriObject::EScene) {                 // <scene> carries no type attribute
                    checkAttributes(n, {});
                    type = "scene";
                } else {
                    checkAttributes(n, {"type"});
                    type = attr(n, "type");
                }
                result = NoriObjectFactory::createInstance(type, props);
                if ((int) result->getClassType() != tag)
                    throw NoriException("Unexpectedly constructed an object of type <%s> (expected type <%s>): %s",
                                        NoriObject::classTypeName(result->getClassType()),
                                        NoriObject::classTypeName((NoriObject::EClassType) tag), result->toString());
                for (NoriObject *ch : children) { result->addChild(ch); ch->setParent(result); }
                result->activate();
            } else {
                setProperty(n, tag, parentList);
            }
        } catch (const NoriException &e) {
            const std::string prefix = "Error while parsing \"";
            if (std::string(e.what()).compare(0, prefix.size(), prefix) == 0) throw;    // already located
            fail(n, e.what());
        }
        return result;
    }
};

} // namespace

NoriObject *loadFromXML(const std::string &filename) {
    const XmlDocument doc = parseXmlFile(filename);
    SceneBuilder b{filename, doc, TransformBuilder()};
    PropertyList list;
    return b.build(doc.root(), list, TInvalid);
}

NORI_NAMESPACE_END
