// node_graph.cpp — headless restatement of the editor graph and its memoised evaluator
// (src/csg_node_graph.rs:235-309).  The egui widgets/drawing of that file are UI and out of scope.
#include <map>
#include <set>
#include <stdexcept>

#include "rm_host.hpp"

namespace ray_marching {

CSGNodeGraph::NodeId CSGNodeGraph::add_root() {
    Node n;
    n.alive = true;
    n.root = true;
    // csg_node_graph.rs:130-138: one connection-only input "SDF" holding CSGNode(None)
    n.inputs.push_back(Input{"SDF", ValueType::CSGNode(std::nullopt), -1});
    nodes_.push_back(std::move(n));
    return (NodeId)nodes_.size() - 1;
}

CSGNodeGraph::NodeId CSGNodeGraph::add_node(const csg::CSGNodeTemplate &t) {
    Node n;
    n.alive = true;
    n.tmpl = t;
    for (auto &p : t.input_params()) n.inputs.push_back(Input{p.first, p.second, -1});  // :141-150
    nodes_.push_back(std::move(n));
    return (NodeId)nodes_.size() - 1;
}

void CSGNodeGraph::remove_node(NodeId id) {
    if (id < 0 || id >= (NodeId)nodes_.size()) throw std::out_of_range("node id");
    nodes_[id].alive = false;
    for (auto &n : nodes_)
        for (auto &in : n.inputs)
            if (in.connection == id) in.connection = -1;
}

void CSGNodeGraph::set_input(NodeId node, const std::string &param, const ValueType &value) {
    if (node < 0 || node >= (NodeId)nodes_.size() || !nodes_[node].alive) throw std::out_of_range("node id");
    for (auto &in : nodes_[node].inputs)
        if (in.name == param) {
            in.value = value;
            return;
        }
    throw std::out_of_range("no input param named " + param);
}

void CSGNodeGraph::connect(NodeId from_node, NodeId to_node, const std::string &to_param) {
    if (from_node < 0 || from_node >= (NodeId)nodes_.size() || !nodes_[from_node].alive || nodes_[from_node].root)
        throw std::out_of_range("source node has no SDF output");
    if (to_node < 0 || to_node >= (NodeId)nodes_.size() || !nodes_[to_node].alive) throw std::out_of_range("node id");
    for (auto &in : nodes_[to_node].inputs)
        if (in.name == to_param) {
            // csg_node_graph.rs:17-22,84-101: ports carry a DataType and the editor only joins ports of EQUAL type.  Every
            // node's single output is DataType::CSGNode ("SDF", :153-158), so a Scalar / Vec3 input refuses the connection
            // (left as it was) instead of failing later in to_vec3().unwrap()
            if (in.value.tag != ValueType::Tag::CSGNode)
                throw std::invalid_argument("DataType mismatch: an SDF output cannot feed the Scalar / Vec3 input " + to_param);
            in.connection = from_node;
            return;
        }
    throw std::out_of_range("no input param named " + to_param);
}

void CSGNodeGraph::disconnect(NodeId to_node, const std::string &to_param) {
    if (to_node < 0 || to_node >= (NodeId)nodes_.size()) throw std::out_of_range("node id");
    for (auto &in : nodes_[to_node].inputs)
        if (in.name == to_param) in.connection = -1;
}

// csg_node_graph.rs:264-309
struct Evaluator {
    const CSGNodeGraph &graph;
    std::map<int, std::optional<ValueType>> output_cache;  // one "SDF" output per node
    std::set<int> in_progress;                             // the reference would recurse forever on a cycle

    explicit Evaluator(const CSGNodeGraph &g) : graph(g) {}

    // :277-282 — connected: value of the upstream output, falling back to the input's own constant
    // when that output evaluates to None; unconnected: the constant.
    ValueType evaluate_input(const CSGNodeGraph::Input &in) {
        if (in.connection >= 0) {
            std::optional<ValueType> v = evaluate_output(in.connection);
            if (v) return *v;
        }
        return in.value;
    }

    // :284-289
    std::optional<ValueType> evaluate_output(int node_id) {
        if (!output_cache.count(node_id)) evaluate_node(node_id);
        auto it = output_cache.find(node_id);
        if (it == output_cache.end()) return std::nullopt;
        return it->second;
    }

    // :291-308
    void evaluate_node(int node_id) {
        const auto &node = graph.nodes_[node_id];
        if (!node.alive || node.root) return;  // NodeTemplate::Root => {}
        if (!in_progress.insert(node_id).second) throw std::runtime_error("cycle in CSG node graph");
        std::map<std::string, ValueType> params;
        for (const auto &in : node.inputs) params[in.name] = evaluate_input(in);
        std::optional<csg::CSGNode> out = node.tmpl.evaluate(params);
        ++graph.evals_;
        output_cache[node_id] = out ? std::optional<ValueType>(ValueType::CSGNode(*out)) : std::nullopt;
        in_progress.erase(node_id);
    }
};

std::optional<csg::CSGNode> CSGNodeGraph::evaluate_root() const {
    evals_ = 0;
    for (const auto &n : nodes_) {  // first Root node, :252-257
        if (!n.alive || !n.root) continue;
        Evaluator ev(*this);
        return ev.evaluate_input(n.inputs.at(0)).to_csg_node();  // :258-260
    }
    return std::nullopt;
}

}  // namespace ray_marching
