#pragma once
#include "quill/Logger.h"
// Log lines are informational only; the shim swallows them (arguments are still evaluated).
namespace quill { namespace shim { template <typename... A> inline void sink(A &&...) {} } }
#define LOG_INFO(logger, ...)    ::quill::shim::sink(logger, __VA_ARGS__)
#define LOG_ERROR(logger, ...)   ::quill::shim::sink(logger, __VA_ARGS__)
#define LOG_WARNING(logger, ...) ::quill::shim::sink(logger, __VA_ARGS__)
