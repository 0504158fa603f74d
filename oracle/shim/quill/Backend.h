// TEST INFRASTRUCTURE (oracle shim): no-op stand-in for quill 10.0.1 (subprojects/quill.wrap).
#pragma once
#include <optional>
#include <functional>
#include <cstring>
#include <limits>
#include <string>
#include <memory>
#include <cstdio>
namespace quill {
struct Backend { static void start() {} static void stop() {} };
}
