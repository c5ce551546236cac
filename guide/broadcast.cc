/*!
 * rabit-b200 guide: broadcast a std::string from rank 0 (cf. reference guide/broadcast.cc).
 */
#include <rabit/rabit.h>

#include <cstdio>
#include <string>

int main(int argc, char *argv[]) {
  rabit::Init(argc, argv);
  std::string msg;
  if (rabit::GetRank() == 0) msg = "hello world";
  std::printf("@node[%d] before-broadcast: s=\"%s\"\n", rabit::GetRank(), msg.c_str());
  rabit::Broadcast(&msg, 0);
  std::printf("@node[%d] after-broadcast: s=\"%s\"\n", rabit::GetRank(), msg.c_str());
  rabit::Finalize();
  return 0;
}
