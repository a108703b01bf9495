// table.h -- internal accessors of gs_table for the other translation units
#pragma once
#include <string>

#include "gpushare.h"

namespace gs_internal {
const std::string &table_uuid(const gs_table *t, uint32_t gpu);
int table_find_uuid(const gs_table *t, const char *uuid, size_t len);
}  // namespace gs_internal
