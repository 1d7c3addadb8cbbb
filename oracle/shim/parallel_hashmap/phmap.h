// Build shim (test infrastructure): maps the reference's parallel-hashmap
// container (a fetched dependency, absent offline) onto the std container
// with the same interface. Container only - no arithmetic of the hot path.
#pragma once
#include <unordered_map>
namespace phmap {
template <class K, class V, class H = std::hash<K>, class E = std::equal_to<K>>
using parallel_flat_hash_map = std::unordered_map<K, V, H, E>;
}
