// Replacement for src/libcore/tls.cpp (which needs boost::multi_index): the ThreadLocalBase interface of
// include/mitsuba/core/tls.h on C++11 thread_local storage.  Infrastructure only -- no arithmetic.
#include <mitsuba/core/tls.h>
#include <unordered_map>
#include <mutex>
MTS_NAMESPACE_BEGIN
namespace detail {
namespace {
struct Entry { void *data; void (*destruct)(void *); };
struct PerThread {
    std::unordered_map<const void *, Entry> map;
    ~PerThread() { for (auto &kv : map) if (kv.second.data) kv.second.destruct(kv.second.data); }
};
thread_local PerThread *t_data = nullptr;
PerThread &local() { if (!t_data) t_data = new PerThread(); return *t_data; }
}
struct ThreadLocalBase::ThreadLocalPrivate {
    ConstructFunctor constructFunctor; DestructFunctor destructFunctor;
    ThreadLocalPrivate(const ConstructFunctor &c, const DestructFunctor &d) : constructFunctor(c), destructFunctor(d) {}
    void *get(bool &existed) {
        PerThread &p = local();
        auto it = p.map.find(this);
        if (it != p.map.end()) { existed = true; return it->second.data; }
        existed = false;
        void *data = constructFunctor();
        p.map[this] = Entry{data, destructFunctor};
        return data;
    }
};
ThreadLocalBase::ThreadLocalBase(const ConstructFunctor &c, const DestructFunctor &d) : d(new ThreadLocalPrivate(c, d)) {}
ThreadLocalBase::~ThreadLocalBase() {
    // entries of other threads are released when those threads exit
    PerThread &p = local();
    auto it = p.map.find(d.get());
    if (it != p.map.end()) { if (it->second.data) it->second.destruct(it->second.data); p.map.erase(it); }
}
void *ThreadLocalBase::get() { bool e; return d->get(e); }
const void *ThreadLocalBase::get() const { bool e; return d->get(e); }
void *ThreadLocalBase::get(bool &existed) { return d->get(existed); }
const void *ThreadLocalBase::get(bool &existed) const { return d->get(existed); }
void initializeGlobalTLS() {}
void destroyGlobalTLS() {}
void initializeLocalTLS() { local(); }
void destroyLocalTLS() { delete t_data; t_data = nullptr; }
}
MTS_NAMESPACE_END
