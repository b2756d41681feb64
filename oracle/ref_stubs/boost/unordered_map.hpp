// TEST INFRASTRUCTURE: stand-in for boost::unordered_map<unsigned long, int> as used by pointHashmap
// (hasher.h:98, hasher.cpp:13-23,28-31): find / end / operator[] / at.  Iteration order is never observed by the
// reference, so any correct associative container is equivalent.  Own chained hash table: std::unordered_map must not
// be used because hasher.h says "using namespace std; using namespace boost;" and names the type unqualified.
#ifndef REF_STANDIN_BOOST_UNORDERED_MAP
#define REF_STANDIN_BOOST_UNORDERED_MAP
#include <vector>
#include <stdexcept>
#include <utility>
namespace boost {
template <typename K, typename V> class unordered_map {
    struct Node { std::pair<K, V> kv; int next; };
    std::vector<Node> nodes_;
    std::vector<int> heads_;
    static size_t mix(K k) { unsigned long long x = (unsigned long long)k; x ^= x >> 33; x *= 0xff51afd7ed558ccdULL; x ^= x >> 33; return (size_t)x; }
    void rehash(size_t nb) {
        heads_.assign(nb, -1);
        for (size_t i = 0; i < nodes_.size(); i++) { size_t b = mix(nodes_[i].kv.first) & (nb - 1); nodes_[i].next = heads_[b]; heads_[b] = (int)i; }
    }
    int locate(const K& k) const {
        if (heads_.empty()) return -1;
        for (int i = heads_[mix(k) & (heads_.size() - 1)]; i >= 0; i = nodes_[i].next) if (nodes_[i].kv.first == k) return i;
        return -1;
    }
public:
    typedef const std::pair<K, V>* const_iterator;
    typedef std::pair<K, V>* iterator;
    unordered_map() {}
    iterator end() { return 0; }
    iterator find(const K& k) { int i = locate(k); return i < 0 ? 0 : &nodes_[i].kv; }
    V& operator[](const K& k) {
        int i = locate(k);
        if (i >= 0) return nodes_[i].kv.second;
        if (nodes_.size() + 1 > heads_.size()) rehash(heads_.empty() ? 1024 : heads_.size() * 2);
        Node n; n.kv.first = k; n.kv.second = V();
        size_t b = mix(k) & (heads_.size() - 1);
        n.next = heads_[b]; heads_[b] = (int)nodes_.size();
        nodes_.push_back(n);
        return nodes_.back().kv.second;
    }
    V& at(const K& k) { int i = locate(k); if (i < 0) throw std::out_of_range("unordered_map::at"); return nodes_[i].kv.second; }
    size_t size() const { return nodes_.size(); }
};
}  // namespace boost
#endif
