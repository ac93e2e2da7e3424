// TEST INFRASTRUCTURE (oracle/) -- what src/extern_template.in.cpp generates for TEMPLATE=kcount
// (src/kcount/CMakeLists.txt:3-10).
#include "kcount.hpp"
#include "kcount.cpp"
__MACRO_KCOUNT__(32, template);
__MACRO_KCOUNT__(64, template);
__MACRO_KCOUNT__(96, template);
__MACRO_KCOUNT__(128, template);
