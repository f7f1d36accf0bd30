#pragma once
#include <variant>
namespace boost {
template <class... T> using variant = std::variant<T...>;
template <class R = void> struct static_visitor { typedef R result_type; };
template <class V, class Var> auto apply_visitor(V &&v, Var &&var) { return std::visit(std::forward<V>(v), std::forward<Var>(var)); }
template <class T, class... U> T &get(std::variant<U...> &v) { return std::get<T>(v); }
template <class T, class... U> const T &get(const std::variant<U...> &v) { return std::get<T>(v); }
template <class T, class... U> T *get(std::variant<U...> *v) { return std::get_if<T>(v); }
template <class T, class... U> const T *get(const std::variant<U...> *v) { return std::get_if<T>(v); }
struct bad_get : std::bad_variant_access {};
}
