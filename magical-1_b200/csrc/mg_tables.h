/* mg_tables.h -- accessor for the host copy of the constant geometry tables (mg_tables.cpp). */
#pragma once
#include "mg_core.cuh"
const MgTables *mg_tables_host();
