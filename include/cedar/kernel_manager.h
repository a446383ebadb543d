// API surface after Cedar (c) 2016-2017 Cedar Developers, BSD-3-Clause -- see NOTICE at the repository root.
// cedar/kernel_manager.h -- the registry: add<K,Impl>(name), set<K>(name),
// setup<K>(args...), run<K>(args...), add_halo(ptr).  Reference:
// include/cedar/kernel_manager.h:11-94.  A missing kernel is logged, not thrown.
#ifndef CEDAR_KERNEL_MANAGER_H
#define CEDAR_KERNEL_MANAGER_H

#include <cedar/type_list.h>
#include <cedar/kernel_params.h>
#include <cedar/config.h>
#include <cedar/halo_exchanger_base.h>

namespace cedar {

template <class tlist> class kernel_manager : public type_map<tlist>
{
public:
	using parent = type_map<tlist>;
	using mtype = typename parent::mtype;
	using parent::kerns;

	kernel_manager(std::shared_ptr<kernel_params> params) : params(params) {}
	kernel_manager(config &conf) { params = build_kernel_params(conf); }

	template <class T, class rclass, class... Args> void add(const std::string &name, Args &&...args)
	{
		parent::template add<T, rclass>(name, std::forward<Args>(args)...);
		this->init<T>(name);
	}
	template <class T, class... Args> void setup(Args &&...args)
	{
		auto kern = parent::template get_ptr<T>();
		log::debug << "setup kernel <" << T::name() << ">" << std::endl;
		if (!kern) { log::error << "kernel not found: " << T::name() << std::endl; return; }
		kern->setup(std::forward<Args>(args)...);
	}
	template <class T, class... Args> void run(Args &&...args)
	{
		auto kern = parent::template get_ptr<T>();
		log::debug << "running kernel <" << T::name() << ">" << std::endl;
		if (!kern) { log::error << "kernel not found: " << T::name() << std::endl; return; }
		kern->run(std::forward<Args>(args)...);
	}
	template <class T> void init(const std::string &name) { parent::template get<T>(name).add_params(params); }
	void add_halo(halo_exchanger_base *halof)
	{
		std::apply([halof](auto &...maps) {
			auto each = [halof](auto &m) { for (auto &kv : m) kv.second->add_halo(halof); };
			(each(maps), ...);
		}, kerns);
	}
	std::shared_ptr<kernel_params> get_params() { return params; }
protected:
	std::shared_ptr<kernel_params> params;
};

}
#endif
