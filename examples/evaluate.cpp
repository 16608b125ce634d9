/*
 * evaluate.cpp - the evaluation half of the reference's program (main.cpp:12-35 + what networkSolver::trainNet does every
 * log interval, src/networksolver.cpp:286,307-308,332,377-379) with the B200 mirror in place of Caffe / OpenCV / Eigen:
 *
 *     datasetManager db(config);  db.generateDatasets();
 *     Net testCNN(network_path + net_name + ".prototxt");  [trained weights]
 *     networkEvaluator::saveLog(testCNN, db, config, iter, time);  networkEvaluator::saveConfusionMatrix(testCNN, db, config);
 *
 * usage: evaluate <config.ini> [weights.pew | -] [iter]
 *   config.ini      the reference's file (paths.hdf5_path holds realSamples_/synthSamples_/templates_<model>.pess,
 *                   paths.network_path + train.net_name + ".prototxt" the net, paths.output_path receives the logs)
 *   weights.pew     PEW1 weight file (tools/caffemodel_to_pew.py from a trained .caffemodel); "-" or absent: the
 *                   prototxt's fillers (xavier / constant), i.e. an untrained net
 * Prints the test-set histogram row it appended to log<log_name>.
 */
#include <chrono>
#include <cstdio>

#include "datasetmanager_b200.hpp"

int main(int argc, char* argv[]) {
    if (argc < 2) {
        std::cerr << "Specify config file as argument" << std::endl;      /* main.cpp:16-20 */
        return 0;
    }
    try {
        const std::string config(argv[1]);
        const std::string weights = argc > 2 ? argv[2] : "-";
        const int iter = argc > 3 ? std::atoi(argv[3]) : 0;
        const auto t0 = std::chrono::steady_clock::now();

        sz::datasetManager db(config);
        db.generateDatasets();

        sz::IniFile pt(config);
        const std::string net_path = pt.get<std::string>("paths.network_path") + pt.get<std::string>("train.net_name") + ".prototxt";
        sz::Net testCNN(net_path);
        if (weights == "-") testCNN.fill(20240607, 0);
        else testCNN.loadWeights(weights);

        const std::vector<int> rotInv = sz::to_array<int>(pt.get<std::string>("input.rotInv"));
        const std::vector<float> bins = {-1, 0, 10, 20, 40, 180};
        const std::vector<float> hist = sz::networkEvaluator::computeHistogram(testCNN, db.getTemplateSet(), db.getTestSet(), rotInv, bins,
                                                                               pt.get<int>("output.kNN"));
        const float secs = std::chrono::duration<float>(std::chrono::steady_clock::now() - t0).count();
        sz::networkEvaluator::saveLog(testCNN, db, config, iter, secs);
        sz::networkEvaluator::saveConfusionMatrix(testCNN, db, config);

        std::printf("objects %zu  templates/object %zu  training/object %zu  test/object %zu\n", db.getTemplateSet().size(),
                    db.getTemplateSet()[0].size(), db.getTrainingSet()[0].size(), db.getTestSet()[0].size());
        std::printf("test");
        for (size_t i = 0; i < bins.size(); ++i) std::printf("\t<%g: %.3f%%", bins[i], hist[i] * 100);
        std::printf("\tmean %.4f\tmedian %.4f\n", hist[bins.size()], hist[bins.size() + 1]);
        return 0;
    } catch (const std::exception& e) {
        std::cerr << "evaluate: " << e.what() << std::endl;
        return 1;
    }
}
