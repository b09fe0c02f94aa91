// rt file.svati output.png — see cli.cpp (gpu/rt.cpp:54-97).
#include "rt_host.h"
int main(int argc, char **argv) { return rtb_cli_main(argc, argv); }
