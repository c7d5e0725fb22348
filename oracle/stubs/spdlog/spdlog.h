// Test-infrastructure stub: the reference's spdlog submodule is not vendored
// (/root/reference/vendor/spdlog is empty). The hot path only uses LOG_DEBUG
// (reference src/system_collisions.cpp:154 via include/log.h:12-19), so every
// logging macro compiles away. <functional> is included here because
// system_collisions.cpp uses std::function (:210) without including it; in the
// real build spdlog leaks it in.
#pragma once
#include <functional>
#define SPDLOG_TRACE(...) ((void)0)
#define SPDLOG_DEBUG(...) ((void)0)
#define SPDLOG_INFO(...) ((void)0)
#define SPDLOG_WARN(...) ((void)0)
#define SPDLOG_ERROR(...) ((void)0)
#define SPDLOG_CRITICAL(...) ((void)0)
