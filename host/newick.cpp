// host/newick.cpp -- see newick.hpp.
#include "newick.hpp"

namespace lazac_host {

namespace {

bool is_sep(char c) { return c == '(' || c == ')' || c == ',' || c == ';'; }

struct Reader {
    const std::string& s;
    size_t pos = 0;
    int unnamed = 0;
    NewickTree t;
    explicit Reader(const std::string& text) : s(text) {}

    // the text up to the next separator ("" when a separator is next or the input ended)
    std::string label_here() const {
        size_t e = pos;
        while (e < s.size() && !is_sep(s[e])) e++;
        return s.substr(pos, e - pos);
    }

    // "[^:]*:?(.*)": only the part before the first ':' names the vertex (a length is read and dropped;
    // std::stof failing on it is the reference's std::invalid_argument)
    static std::string name_of(const std::string& label) {
        size_t c = label.find(':');
        if (c != std::string::npos && c + 1 < label.size()) (void)std::stof(label.substr(c + 1));
        return c == std::string::npos ? label : label.substr(0, c);
    }

    void subtree(int node) {
        if (pos >= s.size() || s[pos] != '(') {
            // a leaf: its label, or a stray separator that is stepped over
            std::string label = label_here();
            if (label.empty()) {
                if (pos < s.size()) pos++;
                else t.name[node] = "";  // the reference parses the empty end-of-input token as a name
            } else {
                t.name[node] = name_of(label);
                pos += label.size();
            }
            return;
        }
        pos++;  // '('
        while (true) {
            if (pos < s.size() && s[pos] == '(') {
                int v = t.add("internal_" + std::to_string(unnamed++));
                t.children[node].push_back(v);
                subtree(v);
            } else if (pos < s.size() && s[pos] == ',') {
                pos++;
            } else if (pos < s.size() && s[pos] == ')') {
                pos++;
                break;
            } else if (pos >= s.size()) {
                throw malformed_parse_exception("Expected right parentheses.");
            } else {
                int v = t.add("");
                t.children[node].push_back(v);
                subtree(v);
            }
        }
        std::string label = label_here();
        if (!label.empty()) {
            t.name[node] = name_of(label);
            pos += label.size();
        } else if (pos >= s.size()) {
            t.name[node] = "";  // same end-of-input quirk (a tree text without the closing ';')
        }
    }
};

}  // namespace

NewickTree read_newick(const std::string& text) {
    Reader r(text);
    int root = r.t.add("root");
    r.subtree(root);
    return r.t;
}

std::string write_newick(const std::vector<std::string>& name, const std::vector<std::vector<int>>& children,
                         const std::vector<std::optional<int>>& len) {
    std::string out;
    struct Frame { int u; size_t next; };
    std::vector<Frame> st;
    st.push_back({0, 0});
    while (!st.empty()) {
        const int u = st.back().u;
        const size_t k = st.back().next;
        if (children[u].empty()) {
            out += name[u];
            if (len[u].has_value()) out += ":" + std::to_string(len[u].value());
            st.pop_back();
            continue;
        }
        if (k == 0) out += "(";
        if (k < children[u].size()) {
            if (k != 0) out += ",";
            st.back().next = k + 1;
            st.push_back({children[u][k], 0});
            continue;
        }
        out += ")";
        out += name[u];
        if (len[u].has_value()) out += ":" + std::to_string(len[u].value());
        st.pop_back();
    }
    out += ";";
    return out;
}

}  // namespace lazac_host
