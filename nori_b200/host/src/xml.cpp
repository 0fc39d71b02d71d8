// xml.cpp — recursive-descent reader for the XML subset of xml.h.
#include "xml.h"
#include <cctype>
#include <cstring>
#include <fstream>
#include <sstream>

NORI_NAMESPACE_BEGIN

namespace {

struct Reader {
    const std::string &s; const std::string &file; const XmlDocument &doc;
    size_t p = 0;

    [[noreturn]] void fail(const std::string &what, size_t at) const {
        throw NoriException("Error while parsing \"%s\": %s (at %s)", file, what, doc.where(at));
    }
    bool eof() const { return p >= s.size(); }
    bool startsWith(const char *lit) const { return s.compare(p, strlen(lit), lit) == 0; }
    static bool isSpace(char c) { return c == ' ' || c == '\t' || c == '\n' || c == '\r'; }
    static bool isNameChar(char c) { return isalnum((unsigned char) c) || c == '_' || c == '-' || c == ':' || c == '.'; }
    void skipSpace() { while (!eof() && isSpace(s[p])) ++p; }

    std::string readName() {
        const size_t b = p;
        while (!eof() && isNameChar(s[p])) ++p;
        if (p == b) fail("expected a name", b);
        return s.substr(b, p - b);
    }

    std::string decode(const std::string &raw, size_t at) const {
        std::string out; out.reserve(raw.size());
        for (size_t i = 0; i < raw.size(); ++i) {
            if (raw[i] != '&') { out += raw[i]; continue; }
            const size_t semi = raw.find(';', i);
            if (semi == std::string::npos) fail("unterminated entity", at);
            const std::string e = raw.substr(i + 1, semi - i - 1);
            if (e == "lt") out += '<'; else if (e == "gt") out += '>'; else if (e == "amp") out += '&';
            else if (e == "quot") out += '"'; else if (e == "apos") out += '\'';
            else if (!e.empty() && e[0] == '#') {
                const long code = e.size() > 1 && (e[1] == 'x' || e[1] == 'X') ? strtol(e.c_str() + 2, nullptr, 16) : strtol(e.c_str() + 1, nullptr, 10);
                if (code <= 0 || code > 127) fail("unsupported character reference", at);
                out += (char) code;
            } else fail("unknown entity \"&" + e + ";\"", at);
            i = semi;
        }
        return out;
    }

    void readAttributes(XmlNode &n) {
        for (;;) {
            skipSpace();
            if (eof()) fail("unexpected end of file inside a tag", n.offset);
            if (s[p] == '>' || s[p] == '/' || s[p] == '?') return;
            const size_t at = p;
            const std::string key = readName();
            skipSpace();
            if (eof() || s[p] != '=') fail("expected '=' after attribute \"" + key + "\"", at);
            ++p; skipSpace();
            if (eof() || (s[p] != '"' && s[p] != '\'')) fail("attribute value must be quoted", at);
            const char q = s[p++];
            const size_t b = p;
            while (!eof() && s[p] != q) ++p;
            if (eof()) fail("unterminated attribute value", at);
            n.attributes.emplace_back(key, decode(s.substr(b, p - b), at));
            ++p;
        }
    }

    /// Reads nodes until the closing tag of `parentName` (or end of file at top level).
    void readContent(std::vector<XmlNode> &out, const std::string *parentName, size_t parentAt) {
        for (;;) {
            const size_t textStart = p;
            while (!eof() && s[p] != '<') ++p;
            if (p > textStart) {
                const std::string raw = s.substr(textStart, p - textStart);
                bool blank = true;
                for (char c : raw) if (!isSpace(c)) { blank = false; break; }
                if (!blank) { XmlNode t; t.kind = XmlNode::Text; t.name = raw; t.offset = textStart; out.push_back(t); }
            }
            if (eof()) {
                if (parentName) fail("missing closing tag </" + *parentName + ">", parentAt);
                return;
            }
            const size_t at = p;
            if (startsWith("<!--")) {
                const size_t e = s.find("-->", p + 4);
                if (e == std::string::npos) fail("unterminated comment", at);
                XmlNode c; c.kind = XmlNode::Comment; c.offset = at; out.push_back(c);
                p = e + 3;
            } else if (startsWith("<?")) {
                const size_t e = s.find("?>", p + 2);
                if (e == std::string::npos) fail("unterminated declaration", at);
                XmlNode d; d.kind = XmlNode::Declaration; d.offset = at; out.push_back(d);
                p = e + 2;
            } else if (startsWith("<!")) {                       // DOCTYPE and friends: skipped
                const size_t e = s.find('>', p);
                if (e == std::string::npos) fail("unterminated markup declaration", at);
                p = e + 1;
            } else if (startsWith("</")) {
                p += 2;
                const std::string name = readName();
                skipSpace();
                if (eof() || s[p] != '>') fail("malformed closing tag", at);
                ++p;
                if (!parentName || name != *parentName) fail("unexpected closing tag </" + name + ">", at);
                return;
            } else {
                ++p;
                XmlNode n; n.kind = XmlNode::Element; n.offset = at;
                n.name = readName();
                readAttributes(n);
                if (s[p] == '/') {
                    if (p + 1 >= s.size() || s[p + 1] != '>') fail("malformed tag", at);
                    p += 2;
                } else if (s[p] == '>') {
                    ++p;
                    readContent(n.children, &n.name, at);
                } else fail("malformed tag", at);
                out.push_back(std::move(n));
            }
        }
    }
};

} // namespace

std::string XmlDocument::where(size_t offset) const {
    size_t line = 1, lineStart = 0;
    for (size_t i = 0; i < offset && i < text.size(); ++i)
        if (text[i] == '\n') { ++line; lineStart = i + 1; }
    char buf[64];
    snprintf(buf, sizeof(buf), "line %zu, col %zu", line, offset - lineStart + 1);
    return buf;
}

const XmlNode &XmlDocument::root() const {
    for (const XmlNode &n : nodes) if (n.kind == XmlNode::Element) return n;
    throw NoriException("XML document has no root element");
}

XmlDocument parseXmlFile(const std::string &filename) {
    std::ifstream is(filename, std::ios::binary);
    if (is.fail()) throw NoriException("Error while parsing \"%s\": File was not found", filename);
    std::stringstream ss; ss << is.rdbuf();
    XmlDocument doc;
    doc.text = ss.str();
    Reader r{doc.text, filename, doc};
    r.readContent(doc.nodes, nullptr, 0);
    int roots = 0;
    for (const XmlNode &n : doc.nodes) if (n.kind == XmlNode::Element) ++roots;
    if (roots != 1) throw NoriException("Error while parsing \"%s\": expected exactly one root element", filename);
    return doc;
}

NORI_NAMESPACE_END
