// Build shim (test infrastructure): phmap::btree_map -> std::map.
#pragma once
#include <map>
namespace phmap {
template <class K, class V> using btree_map = std::map<K, V>;
}
