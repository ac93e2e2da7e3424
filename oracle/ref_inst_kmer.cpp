// TEST INFRASTRUCTURE (oracle/) -- what src/extern_template.in.cpp generates for TEMPLATE=kmer,
// for every MAX_K the reference builds (src/CMakeLists.txt:204-211): include the .cpp, instantiate.
#include "kmer.hpp"
#include "kmer.cpp"
__MACRO_KMER__(32, template);
__MACRO_KMER__(64, template);
__MACRO_KMER__(96, template);
__MACRO_KMER__(128, template);
