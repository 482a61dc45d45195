// Handle types behind the opaque pointers of include/gsparse_b200.h.
#pragma once
#include "../../include/gsparse_b200.h"
#include "host/operators.hpp"

#include <memory>
#include <string>
#include <vector>

struct gsh_metadata {
    std::shared_ptr<gs::meta_data_set> ms;
};

struct gsh_session {
    std::shared_ptr<gs::meta_data_set> ms;
    std::shared_ptr<gs::code_generator> cg;
    gs::operator_executer ex;
    std::vector<std::string> records;
};

namespace gs {
extern thread_local std::string g_last_error;
gs_status fail(gs_status st, const std::string &msg);
gs_status translate_exception();
} // namespace gs
