// placeholder main, replaced by the real CLI
#include <cstdio>
int main() { std::puts("b200pt CLI: not built yet"); return 0; }
