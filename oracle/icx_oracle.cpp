// placeholder, replaced below
extern "C" int icxo_abi_version(void) { return 1; }
