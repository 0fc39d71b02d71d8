// xml.h — the small XML subset Nori scene files use: declaration, comments, nested elements with
// quoted attributes, self-closing tags, the five predefined entities. (The reference parses with
// pugixml, ext/pugixml; scene files never use more than this.)
#pragma once
#include <nori/common.h>
#include <utility>

NORI_NAMESPACE_BEGIN

struct XmlNode {
    enum Kind { Element, Comment, Declaration, Text };
    Kind kind = Element;
    std::string name;                                            // element name, or the text
    std::vector<std::pair<std::string, std::string>> attributes; // in document order
    std::vector<XmlNode> children;
    size_t offset = 0;                                           // byte offset of the node in the file

    const std::string *attribute(const std::string &key) const {
        for (const auto &a : attributes) if (a.first == key) return &a.second;
        return nullptr;
    }
};

struct XmlDocument {
    std::vector<XmlNode> nodes;                                  // top level: declaration, comments, one root element
    std::string text;                                            // file contents (for line/column reporting)
    /// "line L, col C" of a byte offset
    std::string where(size_t offset) const;
    /// first element at top level; throws if there is none
    const XmlNode &root() const;
};

/// Throws NoriException("Error while parsing \"file\": ... (at line L, col C)") on malformed input.
XmlDocument parseXmlFile(const std::string &filename);

NORI_NAMESPACE_END
