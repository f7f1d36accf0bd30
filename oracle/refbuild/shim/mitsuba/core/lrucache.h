// Shim override of include/mitsuba/core/lrucache.h (found first on the include path): the same LRUCache
// interface on std::map + std::list instead of boost::bimap.  Used only by the `serialized` mesh loader.
#pragma once
#if !defined(__LRUCACHE_H)
#define __LRUCACHE_H
#include <mitsuba/mitsuba.h>
#include <boost/function.hpp>
#include <map>
#include <list>
MTS_NAMESPACE_BEGIN
template <typename K, typename KComp, typename V> struct LRUCache : public Object {
public:
    LRUCache() : m_capacity(1) {}
    LRUCache(size_t capacity, const boost::function<V(const K &)> &generatorFunction,
             const boost::function<void(const V &)> &cleanupFunction = NULL)
        : m_capacity(capacity), m_generatorFunction(generatorFunction), m_cleanupFunction(cleanupFunction) {}
    virtual ~LRUCache() { if (m_cleanupFunction) for (auto &kv : m_map) m_cleanupFunction(kv.second.first); }
    bool isFull() const { return m_map.size() == m_capacity; }
    V get(const K &k, bool &hit) {
        auto it = m_map.find(k);
        if (it == m_map.end()) {
            const V v = m_generatorFunction(k);
            if (m_map.size() == m_capacity) {
                auto old = m_map.find(m_order.front());
                if (m_cleanupFunction) m_cleanupFunction(old->second.first);
                m_map.erase(old); m_order.pop_front();
            }
            m_order.push_back(k);
            m_map.insert(std::make_pair(k, std::make_pair(v, --m_order.end())));
            hit = false;
            return v;
        }
        m_order.splice(m_order.end(), m_order, it->second.second);
        hit = true;
        return it->second.first;
    }
private:
    size_t m_capacity;
    boost::function<V(const K &)> m_generatorFunction;
    boost::function<void(const V &)> m_cleanupFunction;
    std::list<K> m_order;
    std::map<K, std::pair<V, typename std::list<K>::iterator>, KComp> m_map;
};
MTS_NAMESPACE_END
#endif
