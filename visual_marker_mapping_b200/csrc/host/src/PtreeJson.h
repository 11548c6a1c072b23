// PtreeJson.h — reader/writer for the JSON dialect of boost::property_tree (read_json / write_json), which
// is what the reference's three project files are written in (src/DetectionIO.cpp:15,150,
// src/ReconstructionIO.cpp:96,140, src/CameraUtilities.cpp:64):
//   every leaf is a quoted string, doubles carry 17 significant digits, 4-space pretty printing, an array is
//   a node whose children all have empty keys, an empty node prints as "", '/' is escaped as "\/".
// The reader is a normal JSON parser that also accepts bare numbers / true / false / null as leaves.
// Site-sized projects (C3: 4 M tag observations, a marker_detections.json of > 1 GB) do not go through a
// whole-file tree: read_json_sections() reads the file in 1 MiB chunks and hands over one child of a
// top-level node at a time, Writer emits the same bytes as write_json() without building a tree.
#ifndef VMM_PTREE_JSON_H_
#define VMM_PTREE_JSON_H_
#include <cstdio>
#include <functional>
#include <memory>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

namespace vmm_json
{
struct Node
{
    std::string data;                                   // leaf value (text)
    std::vector<std::pair<std::string, Node> > children;   // ordered, keys may repeat / be empty

    bool empty() const { return children.empty(); }
    const Node* find(const std::string& key) const
    {
        for (const auto& c : children)
            if (c.first == key) return &c.second;
        return nullptr;
    }
    const Node& get_child(const std::string& key) const
    {
        const Node* n = find(key);
        if (!n) throw std::runtime_error("No such node (" + key + ")");
        return *n;
    }
    std::string get_string(const std::string& key) const { return get_child(key).data; }
    double get_double(const std::string& key) const { return get_child(key).value_double(); }
    int get_int(const std::string& key) const { return get_child(key).value_int(); }
    double value_double() const;
    int value_int() const;

    Node& put(const std::string& key, const std::string& value)
    {
        Node n;
        n.data = value;
        children.emplace_back(key, std::move(n));
        return children.back().second;
    }
    Node& put(const std::string& key, double value) { return put(key, format_double(value)); }
    Node& put(const std::string& key, int value) { return put(key, std::to_string(value)); }
    Node& put(const std::string& key, long long value) { return put(key, std::to_string(value)); }
    Node& add_child(const std::string& key, Node child)
    {
        children.emplace_back(key, std::move(child));
        return children.back().second;
    }
    void push_back(Node child) { children.emplace_back(std::string(), std::move(child)); }

    static std::string format_double(double v);   // 17 significant digits, shortest "%.17g" form
    static Node leaf(double v)
    {
        Node n;
        n.data = format_double(v);
        return n;
    }
};

Node read_json(const std::string& filename);          // throws std::runtime_error with file and offset
Node parse_json(const std::string& text, const std::string& what = "<string>");
void write_json(const std::string& filename, const Node& root);
std::string to_json(const Node& root);

// Streaming reader.  The root must be an object; for every top-level key ("section") each child of its value
// (array element or object member) is parsed into a small tree and passed to `element(section, key, node)`
// as soon as it is complete, then dropped.  Returns the sections in file order (a section whose value is a
// leaf such as "" has no elements) so that callers can report "No such node (<key>)" like ptree::get_child.
typedef std::function<void(const std::string& section, const std::string& key, Node& element)> ElementFn;
std::vector<std::string> read_json_sections(const std::string& filename, const ElementFn& element);
std::vector<std::string> parse_json_sections(const std::string& text, const ElementFn& element, const std::string& what = "<string>");

// Streaming writer: byte for byte what write_json() prints for the equivalent tree (4-space indent, quoted
// leaves, a container without children prints as "", '/' escaped).  Children of an array carry no keys.
class Writer
{
public:
    explicit Writer(const std::string& filename);   // the root object is opened by the constructor
    ~Writer();
    void begin_object(const std::string& key = std::string()) { begin(key, false); }
    void begin_array(const std::string& key = std::string()) { begin(key, true); }
    void end();
    void leaf(const std::string& key, const std::string& value);
    void leaf(const std::string& key, double value);
    void leaf(const std::string& key, int value) { leaf(key, std::to_string(value)); }
    void value(double v) { leaf(std::string(), v); }   // array item
    void node(const std::string& key, const Node& subtree);
    void close();   // closes the root, flushes; throws std::runtime_error on a write error

private:
    struct Frame
    {
        bool is_array;
        size_t count;
    };
    void begin(const std::string& key, bool is_array);
    void child_prefix(const std::string& key);
    void flush();
    std::string filename_;
    std::FILE* file_;
    std::string buf_;
    std::vector<Frame> stack_;
};
}
#endif
