// tables.h -- table selector shared between runtime.cu and the GEMM dispatch.
#pragma once
namespace b200 {
int select_table(int E, int M, int withComp, int dnsmp, int allow_unpinned, const signed char** out);
int select_table_v5(int E, int M, int withComp, int dnsmp, const signed char** out);
// Interns a caller-provided host table by content; the returned pointer is stable for the process.
const signed char* intern_table(const signed char* host, int M);
}
