// PtreeJson.cpp — see PtreeJson.h.
#include "PtreeJson.h"

#include <cerrno>
#include <charconv>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <sstream>
#include <vector>

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
    // std::from_chars reads what strtod reads for the decimal / exponent forms the files hold (correctly rounded), several
    // times faster; anything it does not take whole (hex floats, "inf", a leading '+', locale forms) goes to strtod
    const char* b = data.data();
    const char* e = b + data.size();
    double v = 0.0;
    const std::from_chars_result r = std::from_chars(b, e, v, std::chars_format::general);
    if (r.ec == std::errc() && r.ptr == e && b != e) return v;
    errno = 0;
    char* end = nullptr;
    v = std::strtod(data.c_str(), &end);
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
// Input: a whole string, or a file read in 1 MiB chunks (nothing but the current chunk is kept).
struct Input
{
    const char* cur = nullptr;
    const char* lim = nullptr;
    std::FILE* file = nullptr;
    std::vector<char> chunk;
    size_t line = 1;

    bool fill()
    {
        if (!file) return false;
        if (chunk.empty()) chunk.resize((size_t)1 << 20);
        const size_t n = std::fread(chunk.data(), 1, chunk.size(), file);
        cur = chunk.data();
        lim = cur + n;
        return n > 0;
    }
    bool eof() { return cur == lim && !fill(); }
    char peek() { return *cur; }   // only after !eof()
    char get()
    {
        const char c = *cur++;
        if (c == '\n') ++line;
        return c;
    }
};

struct Parser
{
    Input& in;
    const std::string& what;

    [[noreturn]] void fail(const std::string& msg) const { throw std::runtime_error(what + "(" + std::to_string(in.line) + "): " + msg); }
    void ws()
    {
        while (!in.eof()) {
            const char* p = in.cur;
            while (p != in.lim && (*p == ' ' || *p == '\t' || *p == '\r' || *p == '\n')) {
                if (*p == '\n') ++in.line;
                ++p;
            }
            in.cur = p;
            if (p != in.lim) break;
        }
    }
    bool next_is(char c) { return !in.eof() && in.peek() == c; }
    static void utf8(std::string& out, unsigned cp)
    {
        if (cp < 0x80) out += (char)cp;
        else if (cp < 0x800) { out += (char)(0xC0 | (cp >> 6)); out += (char)(0x80 | (cp & 0x3F)); }
        else { out += (char)(0xE0 | (cp >> 12)); out += (char)(0x80 | ((cp >> 6) & 0x3F)); out += (char)(0x80 | (cp & 0x3F)); }
    }
    void string(std::string& out)
    {
        if (!next_is('"')) fail("expected string");
        in.get();
        out.clear();
        while (true) {
            if (in.eof()) fail("unterminated string");
            {   // plain run up to the next quote / backslash inside the current chunk
                const char* p = in.cur;
                while (p != in.lim && *p != '"' && *p != '\\') {
                    if (*p == '\n') ++in.line;
                    ++p;
                }
                out.append(in.cur, (size_t)(p - in.cur));
                in.cur = p;
                if (p == in.lim) continue;
            }
            const char c = in.get();
            if (c == '"') break;
            if (in.eof()) fail("unterminated escape");
            const char e = in.get();
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
                char hex[5] = { 0, 0, 0, 0, 0 };
                for (int i = 0; i < 4; ++i) {
                    if (in.eof()) fail("bad \\u escape");
                    hex[i] = in.get();
                }
                utf8(out, (unsigned)std::strtoul(hex, nullptr, 16));
                break;
            }
            default: fail("invalid escape sequence");
            }
        }
    }
    // bare literal: number, true, false, null — kept as text, like read_json does
    void literal(std::string& out)
    {
        out.clear();
        while (!in.eof()) {
            const char* p = in.cur;
            while (p != in.lim && *p != ',' && *p != '}' && *p != ']' && *p != ' ' && *p != '\n' && *p != '\r' && *p != '\t') ++p;
            out.append(in.cur, (size_t)(p - in.cur));
            in.cur = p;
            if (p != in.lim) break;
        }
        if (out.empty()) fail("expected value");
    }
    // Children of the container that starts at the cursor ('{' or '['); each (key, child) goes to `sink`.
    template <class Sink> void container(const Sink& sink)
    {
        const bool object = in.get() == '{';
        const char close = object ? '}' : ']';
        ws();
        if (next_is(close)) { in.get(); return; }
        std::string key;
        while (true) {
            if (object) {
                ws();
                string(key);
                ws();
                if (!next_is(':')) fail("expected ':'");
                in.get();
            }
            Node child;
            value(child);
            sink(key, child);
            ws();
            if (next_is(',')) { in.get(); continue; }
            if (next_is(close)) { in.get(); break; }
            fail(object ? "expected ',' or '}'" : "expected ',' or ']'");
        }
    }
    void value(Node& n)
    {
        ws();
        if (in.eof()) fail("unexpected end of input");
        const char c = in.peek();
        if (c == '{' || c == '[')
            container([&n](std::string& key, Node& child) {
                if (n.children.empty()) n.children.reserve(4);   // most containers of the project files hold 2-4 children
                n.children.emplace_back(std::move(key), std::move(child));
                key.clear();
            });
        else if (c == '"') string(n.data);
        else literal(n.data);
    }
    void finish()
    {
        ws();
        if (!in.eof()) fail("garbage after data");
    }
    // root object → sections → elements (see read_json_sections in the header)
    std::vector<std::string> sections(const ElementFn& element)
    {
        std::vector<std::string> seen;
        ws();
        if (in.eof()) fail("unexpected end of input");
        if (in.peek() != '{') {   // not an object: nothing to hand out; parsed (and dropped) for its syntax errors only
            Node n;
            value(n);
            finish();
            return seen;
        }
        in.get();
        ws();
        if (next_is('}')) { in.get(); finish(); return seen; }
        std::string section;
        while (true) {
            ws();
            string(section);
            ws();
            if (!next_is(':')) fail("expected ':'");
            in.get();
            ws();
            if (in.eof()) fail("unexpected end of input");
            seen.push_back(section);
            const char c = in.peek();
            if (c == '{' || c == '[') container([&](std::string& key, Node& child) { element(section, key, child); });
            else {
                Node leaf;
                value(leaf);
            }
            ws();
            if (next_is(',')) { in.get(); continue; }
            if (next_is('}')) { in.get(); break; }
            fail("expected ',' or '}'");
        }
        finish();
        return seen;
    }
};

struct FileInput : Input
{
    explicit FileInput(const std::string& filename)
    {
        file = std::fopen(filename.c_str(), "rb");
        if (!file) throw std::runtime_error(filename + ": cannot open file");
    }
    ~FileInput() { std::fclose(file); }
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

void append_escaped(std::string& out, const std::string& in)   // no temporary when nothing needs escaping (keys, numbers)
{
    bool plain = true;
    for (unsigned char c : in)
        if (c < 0x20 || c == '"' || c == '\\' || c == '/') { plain = false; break; }
    if (plain) out += in;
    else out += escape(in);
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
    Input in;
    in.cur = text.data();
    in.lim = text.data() + text.size();
    Parser ps{ in, what };
    Node n;
    ps.value(n);
    ps.finish();
    return n;
}

Node read_json(const std::string& filename)
{
    FileInput in(filename);
    Parser ps{ in, filename };
    Node n;
    ps.value(n);
    ps.finish();
    return n;
}

std::vector<std::string> parse_json_sections(const std::string& text, const ElementFn& element, const std::string& what)
{
    Input in;
    in.cur = text.data();
    in.lim = text.data() + text.size();
    Parser ps{ in, what };
    return ps.sections(element);
}

std::vector<std::string> read_json_sections(const std::string& filename, const ElementFn& element)
{
    FileInput in(filename);
    Parser ps{ in, filename };
    return ps.sections(element);
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

Writer::Writer(const std::string& filename) : filename_(filename), file_(std::fopen(filename.c_str(), "wb"))
{
    if (!file_) throw std::runtime_error(filename + ": cannot open file");
    stack_.push_back(Frame{ false, 0 });
}

Writer::~Writer()
{
    if (file_) std::fclose(file_);
}

void Writer::flush()
{
    if (!buf_.empty() && std::fwrite(buf_.data(), 1, buf_.size(), file_) != buf_.size()) throw std::runtime_error(filename_ + ": write error");
    buf_.clear();
}

// Everything a child prints before its own value: the parent's bracket (first child), the separator, the
// indentation and, inside an object, the key.
void Writer::child_prefix(const std::string& key)
{
    Frame& parent = stack_.back();
    if (parent.count++ == 0) buf_ += parent.is_array ? '[' : '{';
    else buf_ += ',';
    buf_ += '\n';
    buf_.append((size_t)4 * stack_.size(), ' ');
    if (!parent.is_array) {
        buf_ += '"';
        append_escaped(buf_, key);
        buf_ += "\": ";
    }
    if (buf_.size() > ((size_t)1 << 20)) flush();
}

void Writer::begin(const std::string& key, bool is_array)
{
    child_prefix(key);
    stack_.push_back(Frame{ is_array, 0 });
}

void Writer::end()
{
    if (stack_.size() < 2) throw std::logic_error("vmm_json::Writer::end without begin");
    const Frame f = stack_.back();
    stack_.pop_back();
    if (f.count == 0) { buf_ += "\"\""; return; }
    buf_ += '\n';
    buf_.append((size_t)4 * stack_.size(), ' ');
    buf_ += f.is_array ? ']' : '}';
}

void Writer::leaf(const std::string& key, const std::string& value)
{
    child_prefix(key);
    buf_ += '"';
    append_escaped(buf_, value);
    buf_ += '"';
}

void Writer::leaf(const std::string& key, double value)
{
    child_prefix(key);
    // std::to_chars(general, 17) prints what "%.17g" prints (the C++ standard defines it through printf), several times faster
    char text[64];
    text[0] = '"';
    const std::to_chars_result r = std::to_chars(text + 1, text + 60, value, std::chars_format::general, 17);
    *r.ptr = '"';
    buf_.append(text, (size_t)(r.ptr - text) + 1);
}

void Writer::node(const std::string& key, const Node& subtree)
{
    child_prefix(key);
    write_node(buf_, subtree, (int)stack_.size());
}

void Writer::close()
{
    if (stack_.size() != 1) throw std::logic_error("vmm_json::Writer::close with open containers");
    if (stack_.back().count == 0) buf_ += '{';
    buf_ += "\n}\n";
    flush();
    const bool bad = std::fflush(file_) != 0 || std::ferror(file_);
    const bool bad_close = std::fclose(file_) != 0;
    file_ = nullptr;
    if (bad || bad_close) throw std::runtime_error(filename_ + ": write error");
}
}
