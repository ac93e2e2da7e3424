// TEST INFRASTRUCTURE (oracle/) -- dump_kmers (kmer_dht.cpp:274-297) only needs an ostream; write plain text.
#pragma once
#include <fstream>
namespace zstr {
using ofstream = std::ofstream;
}
