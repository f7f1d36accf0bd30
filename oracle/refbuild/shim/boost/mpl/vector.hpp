// Just enough of Boost.MPL for src/libcore/fmtconv.cpp:1171-1198 (type-pair list of the format converters).
#pragma once
namespace boost { namespace mpl {
template <class... T> struct vector { typedef vector type; };
template <class A, class B> struct pair { typedef A first; typedef B second; };
struct _1; struct _2;
template <class E> struct lambda { typedef lambda type; };
template <class S, class T> struct push_back;                       // used only as a placeholder expression
template <class V, class T> struct push_back_c;
template <class... Ts, class T> struct push_back_c<vector<Ts...>, T> { typedef vector<Ts..., T> type; };
// binary metafunction application for the two placeholder expressions that occur
template <class Op, class S, class U> struct apply2;
template <class T, class S, class U> struct apply2<push_back<_1, pair<T, _2> >, S, U> { typedef typename push_back_c<S, pair<T, U> >::type type; };
template <template <class, class> class F, class S, class U> struct apply2<lambda<F<_2, _1> >, S, U> { typedef typename F<U, S>::type type; };
template <class Seq, class State, class Op> struct fold;
template <class State, class Op> struct fold<vector<>, State, Op> { typedef State type; };
template <class T0, class... Ts, class State, class Op> struct fold<vector<T0, Ts...>, State, Op>
    : fold<vector<Ts...>, typename apply2<Op, State, T0>::type, Op> {};
template <class Seq, class Op> struct transform;
template <class... Ts, template <class> class F> struct transform<vector<Ts...>, lambda<F<_1> > > { typedef vector<F<Ts>...> type; };
template <class Seq> struct for_each_impl;
template <class... Ts> struct for_each_impl<vector<Ts...> > { template <class Fn> static void run(Fn f) { (f(Ts()), ...); } };
template <class Seq, class Fn> void for_each(Fn f) { for_each_impl<typename Seq::type>::run(f); }
} }
