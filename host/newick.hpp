// host/newick.hpp -- Newick reader/writer with the reference's node numbering and naming
// (/root/reference/src/tree_io.cxx:57-121, include/tree_io.hpp:62-93).
#pragma once
#include <optional>
#include <stdexcept>
#include <string>
#include <vector>

namespace lazac_host {

struct malformed_parse_exception : std::runtime_error {
    explicit malformed_parse_exception(const std::string& reason)
        : std::runtime_error("Malformed input - failed to parse.\nReason: " + reason) {}
};

// A rooted tree as the reference's digraph<newick_vertex_data> holds it: node ids are assigned in
// the order vertices appear in the text (root = 0 named "root", unnamed internal nodes
// "internal_<k>" counted in the same order), children kept in ascending id order (std::set).
struct NewickTree {
    std::vector<std::string> name;
    std::vector<std::vector<int>> children;
    int add(const std::string& nm) {
        name.push_back(nm);
        children.emplace_back();
        return (int)name.size() - 1;
    }
    int size() const { return (int)name.size(); }
};

NewickTree read_newick(const std::string& text);

// print_newick_tree(tree, 0) + ";" : children in ascending id order, ":<int>" branch lengths where present.
std::string write_newick(const std::vector<std::string>& name, const std::vector<std::vector<int>>& children,
                         const std::vector<std::optional<int>>& in_branch_length);

}  // namespace lazac_host
