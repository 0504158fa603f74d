#pragma once
#include "quill/Backend.h"
namespace quill {
struct Logger {};
struct ConsoleSink {};
struct Frontend {
  template <typename S> static std::shared_ptr<S> create_or_get_sink(const char *) { return std::make_shared<S>(); }
  template <typename S> static Logger *create_or_get_logger(const char *, S &&) { static Logger l; return &l; }
};
}
