// oracle/stubs/boost/intrusive/list.hpp -- TEST INFRASTRUCTURE ONLY.
//
// hbt/src/tagstack/Slicer.h:12 includes <boost/intrusive/list.hpp>; boost is not in this image.
// This is an independent, minimal stand-in for the part Slicer.h uses (list_base_hook with
// link_mode<auto_unlink>, list<T, constant_time_size<false>>: push_front, iteration, clear,
// hook.is_linked()/unlink()), written against the documented behaviour of those classes:
//   * a hook is a node of a circular doubly linked list whose sentinel lives in the list;
//   * copying or moving a hook yields an UNLINKED hook (boost: "the copy is default
//     initialised"), destruction of an auto_unlink hook takes it out of its list;
//   * list::clear() leaves every former member unlinked; the list owns nothing.
#pragma once
#include <cstddef>
#include <iterator>

namespace boost {
namespace intrusive {

struct auto_unlink {};
template <class Mode>
struct link_mode {};
template <bool B>
struct constant_time_size {};

template <class... Options>
class list_base_hook {
 public:
  list_base_hook() noexcept {}
  list_base_hook(const list_base_hook&) noexcept {}
  list_base_hook& operator=(const list_base_hook&) noexcept { return *this; }
  ~list_base_hook() { unlink(); }
  bool is_linked() const noexcept { return next_ != nullptr; }
  void unlink() noexcept {
    if (next_ != nullptr) {
      prev_->next_ = next_;
      next_->prev_ = prev_;
      next_ = prev_ = nullptr;
    }
  }
  list_base_hook* next_ = nullptr;
  list_base_hook* prev_ = nullptr;
};

template <class T, class... Options>
class list {
  using hook = list_base_hook<link_mode<auto_unlink>>;

 public:
  list() noexcept { root_.next_ = root_.prev_ = &root_; }
  list(const list&) = delete;
  list& operator=(const list&) = delete;
  ~list() {
    clear();
    root_.next_ = root_.prev_ = nullptr;  // so that the sentinel's own destructor is a no-op
  }

  class iterator {
   public:
    using iterator_category = std::forward_iterator_tag;
    using value_type = T;
    using difference_type = std::ptrdiff_t;
    using pointer = T*;
    using reference = T&;
    explicit iterator(hook* n) : n_{n} {}
    T& operator*() const { return *static_cast<T*>(n_); }
    T* operator->() const { return static_cast<T*>(n_); }
    iterator& operator++() {
      n_ = n_->next_;
      return *this;
    }
    bool operator==(const iterator& o) const { return n_ == o.n_; }
    bool operator!=(const iterator& o) const { return n_ != o.n_; }

   private:
    hook* n_;
  };

  iterator begin() { return iterator{root_.next_}; }
  iterator end() { return iterator{&root_}; }
  bool empty() const { return root_.next_ == &root_; }

  void push_front(T& v) {
    hook* h = static_cast<hook*>(&v);
    h->unlink();
    h->next_ = root_.next_;
    h->prev_ = &root_;
    root_.next_->prev_ = h;
    root_.next_ = h;
  }

  void clear() {
    hook* n = root_.next_;
    while (n != &root_) {
      hook* nx = n->next_;
      n->next_ = n->prev_ = nullptr;
      n = nx;
    }
    root_.next_ = root_.prev_ = &root_;
  }

 private:
  hook root_;
};

}  // namespace intrusive
}  // namespace boost
