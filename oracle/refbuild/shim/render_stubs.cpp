// src/librender/scenehandler.cpp is the Xerces XML front end and is not part of the oracle build; the one free
// function other TUs use from it (scenehandler.cpp:269, called by the `serialized` mesh cache) lives here.
#include <mitsuba/mitsuba.h>
#include <vector>
MTS_NAMESPACE_BEGIN
typedef void (*CleanupFun)();
static std::vector<CleanupFun> g_cleanup;
extern MTS_EXPORT_RENDER void pushSceneCleanupHandler(void (*cleanup)());
extern MTS_EXPORT_RENDER void runSceneCleanupHandlers();
void pushSceneCleanupHandler(void (*cleanup)()) { g_cleanup.push_back(cleanup); }
void runSceneCleanupHandlers() { for (CleanupFun f : g_cleanup) f(); g_cleanup.clear(); }
MTS_NAMESPACE_END
