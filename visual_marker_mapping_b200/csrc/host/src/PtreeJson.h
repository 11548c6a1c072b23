// PtreeJson.h — reader/writer for the JSON dialect of boost::property_tree (read_json / write_json), which
// is what the reference's three project files are written in (src/DetectionIO.cpp:15,150,
// src/ReconstructionIO.cpp:96,140, src/CameraUtilities.cpp:64):
//   every leaf is a quoted string, doubles carry 17 significant digits, 4-space pretty printing, an array is
//   a node whose children all have empty keys, an empty node prints as "", '/' is escaped as "\/".
// The reader is a normal JSON parser that also accepts bare numbers / true / false / null as leaves.
#ifndef VMM_PTREE_JSON_H_
#define VMM_PTREE_JSON_H_
#include <cstdio>
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
}
#endif
