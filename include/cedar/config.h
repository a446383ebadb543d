// API surface after Cedar (c) 2016-2017 Cedar Developers, BSD-3-Clause -- see NOTICE at the repository root.
// cedar/config.h -- JSON configuration with dotted-path access, the interface
// of the reference include/cedar/config.h (get / get-with-default / getvec /
// set / getconf) over a small built-in JSON reader (no third-party parser).
#ifndef CONFIG_READER_H
#define CONFIG_READER_H

#include <map>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

namespace cedar {

class jvalue
{
public:
	enum class kind { null, boolean, number, string, array, object };
	kind k = kind::null;
	bool b = false;
	double num = 0;
	std::string str;
	std::vector<jvalue> arr;
	std::map<std::string, jvalue> obj;

	static jvalue parse(const std::string &text);   // throws std::runtime_error
	const jvalue *find(const std::string &dotted) const;
	jvalue &at_path(const std::string &dotted);      // creates objects on the way
};

namespace detail {
template <class T> struct jconv;
template <> struct jconv<int> { static int get(const jvalue &v) { if (v.k != jvalue::kind::number) throw std::runtime_error("number expected"); return static_cast<int>(v.num); } static void put(jvalue &v, int x) { v = jvalue(); v.k = jvalue::kind::number; v.num = x; } };
template <> struct jconv<unsigned int> { static unsigned int get(const jvalue &v) { if (v.k != jvalue::kind::number) throw std::runtime_error("number expected"); return static_cast<unsigned int>(v.num); } static void put(jvalue &v, unsigned int x) { v = jvalue(); v.k = jvalue::kind::number; v.num = x; } };
template <> struct jconv<double> { static double get(const jvalue &v) { if (v.k != jvalue::kind::number) throw std::runtime_error("number expected"); return v.num; } static void put(jvalue &v, double x) { v = jvalue(); v.k = jvalue::kind::number; v.num = x; } };
template <> struct jconv<float> { static float get(const jvalue &v) { return static_cast<float>(jconv<double>::get(v)); } static void put(jvalue &v, float x) { jconv<double>::put(v, x); } };
template <> struct jconv<bool> { static bool get(const jvalue &v) { if (v.k != jvalue::kind::boolean) throw std::runtime_error("boolean expected"); return v.b; } static void put(jvalue &v, bool x) { v = jvalue(); v.k = jvalue::kind::boolean; v.b = x; } };
template <> struct jconv<std::string> { static std::string get(const jvalue &v) { if (v.k != jvalue::kind::string) throw std::runtime_error("string expected"); return v.str; } static void put(jvalue &v, const std::string &x) { v = jvalue(); v.k = jvalue::kind::string; v.str = x; } };
template <> struct jconv<const char *> { static void put(jvalue &v, const char *x) { jconv<std::string>::put(v, x); } };
template <class T> struct jconv<std::vector<T>> {
	static std::vector<T> get(const jvalue &v)
	{
		if (v.k != jvalue::kind::array) throw std::runtime_error("array expected");
		std::vector<T> r;
		for (auto &e : v.arr) r.push_back(jconv<T>::get(e));
		return r;
	}
	static void put(jvalue &v, const std::vector<T> &x)
	{
		v = jvalue(); v.k = jvalue::kind::array;
		for (const auto &e : x) { jvalue c; jconv<T>::put(c, e); v.arr.push_back(c); }
	}
};
}

class config
{
public:
	config();                                   // reads ./config.json
	explicit config(const std::string &fname);  // "" = empty configuration
	explicit config(const char *fname) : config(std::string(fname)) {}
	explicit config(jvalue &&tree) : root(std::move(tree)) {}

	template <typename T> void set(std::string path, T val) { detail::jconv<typename std::decay<T>::type>::put(root.at_path(path), val); }
	template <typename T> void setvec(std::string path, std::vector<T> vec) { set<std::vector<T>>(path, vec); }
	template <typename T> T get(std::string path)
	{
		const jvalue *v = root.find(path);
		if (!v) throw std::runtime_error("config: missing key " + path);
		return detail::jconv<T>::get(*v);
	}
	template <typename T> T get(std::string path, T default_value)
	{
		try { return get<T>(path); } catch (...) { return default_value; }
	}
	template <typename T> std::vector<T> getvec(std::string path)
	{
		try { return get<std::vector<T>>(path); } catch (...) { return std::vector<T>(); }
	}
	template <typename T> std::vector<std::vector<T>> getnvec(std::string path) { return get<std::vector<std::vector<T>>>(path); }
	// sub-tree at `path` as its own config (nullptr when absent); "" = copy of the whole tree
	std::shared_ptr<config> getconf(std::string path);
	void set_config(std::string fname);
private:
	void read();
	jvalue root;
	std::string fname;
};

}
#endif
