// Test-infrastructure shim (oracle/_ref build only): a minimal stand-in for
// Boost.Iterator's iterator_facade so that the reference's libbio headers
// compile under g++ in a container without Boost. Not product code.
#pragma once
#include <cstddef>
#include <iterator>
#include <type_traits>
namespace boost {
	struct bidirectional_traversal_tag {};
	struct forward_traversal_tag {};
	struct random_access_traversal_tag {};
	class iterator_core_access {
	public:
		template <typename I> static auto dereference(I const &it) -> decltype(it.dereference()) { return it.dereference(); }
		template <typename I> static void increment(I &it) { it.increment(); }
		template <typename I> static void decrement(I &it) { it.decrement(); }
		template <typename I, typename J> static bool equal(I const &a, J const &b) { return a.equal(b); }
		template <typename I> static void advance(I &it, std::ptrdiff_t n) { it.advance(n); }
		template <typename I, typename J> static std::ptrdiff_t distance_to(I const &a, J const &b) { return a.distance_to(b); }
	};
	template <typename Derived, typename Value, typename Tag, typename Reference = Value &, typename Difference = std::ptrdiff_t>
	class iterator_facade {
		Derived &self() { return *static_cast <Derived *>(this); }
		Derived const &self() const { return *static_cast <Derived const *>(this); }
	public:
		typedef std::remove_const_t <Value> value_type;
		typedef Reference reference;
		typedef Value *pointer;
		typedef Difference difference_type;
		typedef std::bidirectional_iterator_tag iterator_category;
		reference operator*() const { return iterator_core_access::dereference(self()); }
		// Boost returns a proxy when Reference is not a true reference; do the same.
		struct arrow_proxy { std::remove_reference_t <Reference> value; std::remove_reference_t <Reference> *operator->() { return &value; } };
		template <typename R = Reference, std::enable_if_t <std::is_reference_v <R>, int> = 0>
		std::remove_reference_t <R> *operator->() const { return &iterator_core_access::dereference(self()); }
		template <typename R = Reference, std::enable_if_t <!std::is_reference_v <R>, int> = 0>
		arrow_proxy operator->() const { return arrow_proxy{iterator_core_access::dereference(self())}; }
		Derived &operator++() { iterator_core_access::increment(self()); return self(); }
		Derived operator++(int) { Derived tmp(self()); iterator_core_access::increment(self()); return tmp; }
		Derived &operator--() { iterator_core_access::decrement(self()); return self(); }
		Derived operator--(int) { Derived tmp(self()); iterator_core_access::decrement(self()); return tmp; }
		template <typename O> bool operator==(O const &o) const { return iterator_core_access::equal(self(), o); }
		template <typename O> bool operator!=(O const &o) const { return !iterator_core_access::equal(self(), o); }
	};
}
