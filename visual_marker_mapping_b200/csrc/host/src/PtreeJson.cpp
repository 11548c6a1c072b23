// PtreeJson.cpp — see PtreeJson.h.
#include "PtreeJson.h"

#include <cerrno>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <sstream>

namespace vmm_json
{
std::string Node::format_double(double v)
{
    char buf[64];
    std::snprintf(buf, sizeof(buf), "%.17g", v);   // ostream << setprecision(17): what ptree::put(double) does
    return buf;
}

double Node::value_double() const
{
    errno = 0;
    char* end = nullptr;
    const double v = std::strtod(data.c_str(), &end);
    while (end && *end == ' ') ++end;
    if (data.empty() || !end || *end != '\0') throw std::runtime_error("conversion of data to type \"double\" failed: \"" + data + "\"");
    return v;
}

int Node::value_int() const
{
    char* end = nullptr;
    const long v = std::strtol(data.c_str(), &end, 10);
    while (end && *end == ' ') ++end;
    if (data.empty() || !end || *end != '\0') throw std::runtime_error("conversion of data to type \"int\" failed: \"" + data + "\"");
    return (int)v;
}

namespace
{
struct Parser
{
    const std::string& s;
    const std::string& what;
    size_t p = 0;

    [[noreturn]] void fail(const std::string& msg) const
    {
        size_t line = 1;
        for (size_t i = 0; i < p && i < s.size(); ++i)
            if (s[i] == '\n') ++line;
        throw std::runtime_error(what + "(" + std::to_string(line) + "): " + msg);
    }
    void ws()
    {
        while (p < s.size() && (s[p] == ' ' || s[p] == '\t' || s[p] == '\n' || s[p] == '\r')) ++p;
    }
    static void utf8(std::string& out, unsigned cp)
    {
        if (cp < 0x80) out += (char)cp;
        else if (cp < 0x800) { out += (char)(0xC0 | (cp >> 6)); out += (char)(0x80 | (cp & 0x3F)); }
        else { out += (char)(0xE0 | (cp >> 12)); out += (char)(0x80 | ((cp >> 6) & 0x3F)); out += (char)(0x80 | (cp & 0x3F)); }
    }
    std::string string()
    {
        if (s[p] != '"') fail("expected string");
        ++p;
        std::string out;
        while (true) {
            if (p >= s.size()) fail("unterminated string");
            const char c = s[p++];
            if (c == '"') break;
            if (c != '\\') { out += c; continue; }
            if (p >= s.size()) fail("unterminated escape");
            const char e = s[p++];
            switch (e) {
            case '"': out += '"'; break;
            case '\\': out += '\\'; break;
            case '/': out += '/'; break;
            case 'b': out += '\b'; break;
            case 'f': out += '\f'; break;
            case 'n': out += '\n'; break;
            case 'r': out += '\r'; break;
            case 't': out += '\t'; break;
            case 'u': {
                if (p + 4 > s.size()) fail("bad \\u escape");
                utf8(out, (unsigned)std::strtoul(s.substr(p, 4).c_str(), nullptr, 16));
                p += 4;
                break;
            }
            default: fail("invalid escape sequence");
            }
        }
        return out;
    }
    Node value()
    {
        ws();
        if (p >= s.size()) fail("unexpected end of input");
        Node n;
        const char c = s[p];
        if (c == '{') {
            ++p;
            ws();
            if (p < s.size() && s[p] == '}') { ++p; return n; }
            while (true) {
                ws();
                std::string key = string();
                ws();
                if (p >= s.size() || s[p] != ':') fail("expected ':'");
                ++p;
                Node child = value();
                n.children.emplace_back(std::move(key), std::move(child));
                ws();
                if (p < s.size() && s[p] == ',') { ++p; continue; }
                if (p < s.size() && s[p] == '}') { ++p; break; }
                fail("expected ',' or '}'");
            }
        } else if (c == '[') {
            ++p;
            ws();
            if (p < s.size() && s[p] == ']') { ++p; return n; }
            while (true) {
                Node child = value();
                n.children.emplace_back(std::string(), std::move(child));
                ws();
                if (p < s.size() && s[p] == ',') { ++p; continue; }
                if (p < s.size() && s[p] == ']') { ++p; break; }
                fail("expected ',' or ']'");
            }
        } else if (c == '"') {
            n.data = string();
        } else {
            // bare literal: number, true, false, null — kept as text, like read_json does
            const size_t b = p;
            while (p < s.size() && s[p] != ',' && s[p] != '}' && s[p] != ']' && s[p] != ' ' && s[p] != '\n' && s[p] != '\r' && s[p] != '\t') ++p;
            if (p == b) fail("expected value");
            n.data = s.substr(b, p - b);
        }
        return n;
    }
};

std::string escape(const std::string& in)
{
    std::string out;
    out.reserve(in.size());
    for (unsigned char c : in) {
        switch (c) {
        case '"': out += "\\\""; break;
        case '\\': out += "\\\\"; break;
        case '/': out += "\\/"; break;
        case '\b': out += "\\b"; break;
        case '\f': out += "\\f"; break;
        case '\n': out += "\\n"; break;
        case '\r': out += "\\r"; break;
        case '\t': out += "\\t"; break;
        default:
            if (c < 0x20) {
                char buf[8];
                std::snprintf(buf, sizeof(buf), "\\u%04X", c);
                out += buf;
            } else out += (char)c;
        }
    }
    return out;
}

void write_node(std::string& out, const Node& n, int indent)
{
    if (indent > 0 && n.empty()) {
        out += '"';
        out += escape(n.data);
        out += '"';
        return;
    }
    bool is_array = indent > 0;
    if (is_array)
        for (const auto& c : n.children)
            if (!c.first.empty()) { is_array = false; break; }
    out += is_array ? '[' : '{';
    out += '\n';
    for (size_t i = 0; i < n.children.size(); ++i) {
        out.append((size_t)4 * (indent + 1), ' ');
        if (!is_array) {
            out += '"';
            out += escape(n.children[i].first);
            out += "\": ";
        }
        write_node(out, n.children[i].second, indent + 1);
        if (i + 1 != n.children.size()) out += ',';
        out += '\n';
    }
    out.append((size_t)4 * indent, ' ');
    out += is_array ? ']' : '}';
}
}

Node parse_json(const std::string& text, const std::string& what)
{
    Parser ps{ text, what };
    Node n = ps.value();
    ps.ws();
    if (ps.p != text.size()) ps.fail("garbage after data");
    return n;
}

Node read_json(const std::string& filename)
{
    std::ifstream in(filename.c_str(), std::ios::binary);
    if (!in) throw std::runtime_error(filename + ": cannot open file");
    std::stringstream ss;
    ss << in.rdbuf();
    return parse_json(ss.str(), filename);
}

std::string to_json(const Node& root)
{
    std::string out;
    write_node(out, root, 0);
    out += '\n';
    return out;
}

void write_json(const std::string& filename, const Node& root)
{
    std::ofstream out(filename.c_str(), std::ios::binary);
    if (!out) throw std::runtime_error(filename + ": cannot open file");
    const std::string text = to_json(root);
    out.write(text.data(), (std::streamsize)text.size());
    if (!out) throw std::runtime_error(filename + ": write error");
}
}
