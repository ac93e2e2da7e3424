// TEST INFRASTRUCTURE (oracle/) -- nothing from the thread pool is reached on the CPU kcount path.
#pragma once
