// TEST INFRASTRUCTURE (oracle shim) -- not product code.
// Stand-in for ankerl::unordered_dense 4.8.1 (subprojects/unordered_dense.wrap in the
// reference; un-vendored).  Restates the one property the reference's results depend on:
// values live in a dense vector in INSERTION order, iteration walks that vector,
// operator[] on an existing key keeps its slot, erase moves the last element into the hole.
// Used by reference include/types.h:4-16; order-sensitive use: src/evaluate.cpp:913 (`data`).
#pragma once
#include <vector>
#include <unordered_map>
#include <utility>
#include <functional>
#include <cstddef>
#include <cstdint>
#include <string>

namespace ankerl { namespace unordered_dense {

template <typename K, typename V>
class map {
  std::vector<std::pair<K, V>> vals_;
  std::unordered_map<K, size_t> idx_;
public:
  using value_type = std::pair<K, V>;
  using iterator = typename std::vector<value_type>::iterator;
  using const_iterator = typename std::vector<value_type>::const_iterator;
  iterator begin() { return vals_.begin(); }
  iterator end() { return vals_.end(); }
  const_iterator begin() const { return vals_.begin(); }
  const_iterator end() const { return vals_.end(); }
  bool empty() const { return vals_.empty(); }
  size_t size() const { return vals_.size(); }
  void clear() { vals_.clear(); idx_.clear(); }
  void reserve(size_t n) { vals_.reserve(n); idx_.reserve(n); }
  iterator find(const K &k) {
    auto it = idx_.find(k);
    return it == idx_.end() ? vals_.end() : vals_.begin() + it->second;
  }
  const_iterator find(const K &k) const {
    auto it = idx_.find(k);
    return it == idx_.end() ? vals_.end() : vals_.begin() + it->second;
  }
  size_t count(const K &k) const { return idx_.count(k); }
  V &operator[](const K &k) {
    auto it = idx_.find(k);
    if (it != idx_.end()) return vals_[it->second].second;
    idx_.emplace(k, vals_.size());
    vals_.emplace_back(k, V());
    return vals_.back().second;
  }
  std::pair<iterator, bool> insert(const value_type &kv) {
    auto it = idx_.find(kv.first);
    if (it != idx_.end()) return {vals_.begin() + it->second, false};
    idx_.emplace(kv.first, vals_.size());
    vals_.push_back(kv);
    return {vals_.end() - 1, true};
  }
  template <typename... Args>
  std::pair<iterator, bool> emplace(Args &&...args) {
    return insert(value_type(std::forward<Args>(args)...));
  }
  iterator erase(iterator it) {
    size_t pos = it - vals_.begin();
    idx_.erase(it->first);
    if (pos + 1 != vals_.size()) {
      vals_[pos] = std::move(vals_.back());
      idx_[vals_[pos].first] = pos;
    }
    vals_.pop_back();
    return vals_.begin() + pos;
  }
  size_t erase(const K &k) {
    auto it = find(k);
    if (it == end()) return 0;
    erase(it);
    return 1;
  }
};

template <typename K>
class set {
  std::vector<K> vals_;
  std::unordered_map<K, size_t> idx_;
public:
  using iterator = typename std::vector<K>::iterator;
  using const_iterator = typename std::vector<K>::const_iterator;
  iterator begin() { return vals_.begin(); }
  iterator end() { return vals_.end(); }
  const_iterator begin() const { return vals_.begin(); }
  const_iterator end() const { return vals_.end(); }
  bool empty() const { return vals_.empty(); }
  size_t size() const { return vals_.size(); }
  void clear() { vals_.clear(); idx_.clear(); }
  void reserve(size_t n) { vals_.reserve(n); idx_.reserve(n); }
  size_t count(const K &k) const { return idx_.count(k); }
  iterator find(const K &k) {
    auto it = idx_.find(k);
    return it == idx_.end() ? vals_.end() : vals_.begin() + it->second;
  }
  std::pair<iterator, bool> insert(const K &k) {
    auto it = idx_.find(k);
    if (it != idx_.end()) return {vals_.begin() + it->second, false};
    idx_.emplace(k, vals_.size());
    vals_.push_back(k);
    return {vals_.end() - 1, true};
  }
  template <typename... Args>
  std::pair<iterator, bool> emplace(Args &&...args) { return insert(K(std::forward<Args>(args)...)); }
};

}} // namespace
