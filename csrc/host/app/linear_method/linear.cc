// role x {darlin, async_sgd, validation} -> concrete app (ref src/app/linear_method/linear.cc:8-34)
#include "app/linear_method/linear.h"

#include "app/linear_method/async_sgd.h"
#include "app/linear_method/darlin.h"
#include "app/linear_method/model_evaluation.h"

namespace PS {
namespace LM {

App* createApp(const string& name, const ConfigView& conf) {
  Node my_node = Postoffice::current().myNode();
  auto role = my_node.role();
  App* app = nullptr;
  if (conf.has_darlin()) {
    if (role == NodeView::SCHEDULER) app = new DarlinScheduler(name, conf);
    else if (role == NodeView::WORKER) app = new DarlinWorker(name, conf);
    else if (role == NodeView::SERVER) app = new DarlinServer(name, conf);
  } else if (conf.has_async_sgd()) {
    typedef float Real;  // the reference runs online learning in single precision (linear.cc:20)
    if (role == NodeView::SCHEDULER) app = new AsyncSGDScheduler(name, conf);
    else if (role == NodeView::WORKER) app = new AsyncSGDWorker<Real>(name, conf);
    else if (role == NodeView::SERVER) app = new AsyncSGDServer<Real>(name, conf);
  } else if (conf.has_validation_data()) {
    app = new ModelEvaluation(name, conf);
  }
  CHECK(app) << "fail to create " << conf.ShortDebugString() << " at " << my_node.ShortDebugString();
  return app;
}

// the text-format AppConfig shipped by the scheduler -> app
App* createAppFromText(const string& name, const string& conf_text) {
  AppConfig conf;
  string err;
  CHECK(conf.raw()->parse(conf_text, &err)) << "cannot parse the app config: " << err;
  CHECK(conf.has_linear_method()) << "only linear_method apps are built in";
  return createApp(name, conf.linear_method());
}

}  // namespace LM
}  // namespace PS
