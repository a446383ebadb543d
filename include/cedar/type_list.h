// API surface after Cedar (c) 2016-2017 Cedar Developers, BSD-3-Clause -- see NOTICE at the repository root.
// cedar/type_list.h -- type_list / type_cat / type_map: one name->implementation
// map per kernel interface plus the name of the current default.  Same public
// members as the reference (type_list.h:36-124: kerns, defaults, get_ptr, get,
// add, set); the slot lookup is a constexpr search instead of the recursive
// matching_index templates.
#ifndef CEDAR_TYPE_LIST_H
#define CEDAR_TYPE_LIST_H

#include <array>
#include <map>
#include <memory>
#include <string>
#include <tuple>
#include <type_traits>

#include <cedar/util/log.h>

namespace cedar {

template <class... T> struct type_list {};

template <class ta, class tb> struct type_cat;
template <typename... a, typename... b> struct type_cat<type_list<a...>, type_list<b...>>
{
	typedef type_list<a..., b...> type;
};

template <class U> class type_map;

template <typename... types> class type_map<type_list<types...>>
{
public:
	typedef std::tuple<std::map<std::string, std::shared_ptr<types>>...> mtype;
	mtype kerns;
	std::array<std::string, sizeof...(types)> defaults;

	template <class T> static constexpr std::size_t slot()
	{
		constexpr bool hit[] = {std::is_same<T, types>::value...};
		std::size_t i = 0;
		for (; i < sizeof...(types); ++i)
			if (hit[i]) break;
		return i;
	}
	template <class T> std::shared_ptr<T> get_ptr() { return get_ptr<T>(defaults[checked<T>()]); }
	template <class T> std::shared_ptr<T> get_ptr(const std::string &name)
	{
		auto &m = std::get<checked<T>()>(kerns);
		auto it = m.find(name);
		return it == m.end() ? std::shared_ptr<T>() : it->second;
	}
	template <class T> T &get() { return *(get_ptr<T>()); }
	template <class T> T &get(const std::string &name) { return *(get_ptr<T>(name)); }
	template <class T, class rclass, class... Args> void add(const std::string &name, Args &&...args)
	{
		std::get<checked<T>()>(kerns)[name] = std::make_shared<rclass>(std::forward<Args>(args)...);
	}
	template <class T> void set(const std::string &name) { defaults[checked<T>()] = name; }
private:
	template <class T> static constexpr std::size_t checked()
	{
		static_assert(slot<T>() < sizeof...(types), "<type_map> kernel type is not in this registry's type list");
		return slot<T>();
	}
};

}
#endif
