// build/ps: the node binary of the linear-method apps (ref src/app/main/main.cc:15-39)
//   build/ps -my_node <Node> -scheduler <Node> -num_servers S -num_workers W -app_file f
#include "app/linear_method/linear.h"
#include "system/runtime.h"

int main(int argc, char* argv[]) {
  PS::App::setFactory(PS::LM::createAppFromText);
  PS::Postoffice& po = PS::Postoffice::instance();
  po.start(&argc, &argv);
  po.stop();
  return 0;
}
