// TEST INFRASTRUCTURE (oracle/): empty stand-in, see error_handling.h in this directory.
#pragma once
