// TEST INFRASTRUCTURE (oracle/) -- nothing of the thread pool is used by merge_reads.cpp.
#pragma once
