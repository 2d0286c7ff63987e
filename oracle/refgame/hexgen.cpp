// TEST INFRASTRUCTURE (oracle side) -- not product code.
//
// Position generator linked against the UNMODIFIED reference C++ game
// (/root/reference/cpp_hex/hex_graph_game/{graph,hex_board_game,
// shannon_node_switching_game}.cpp, compiled where they lie by oracle/refgame/Makefile).
// It plays seeded uniform-random games with dead/captured removal (the loop of
// cpp_hex/hex_graph_game/tests/speedtest.cpp:33-49) and dumps every non-terminal
// position as the tensors Node_switching_game::convert_graph emits
// (shannon_node_switching_game.cpp:690-726): n, the IS_TERMINAL column and the
// source-major CSR edge list.  Output is text, parsed by tests/golden/make_fixtures.py.
//
// usage: hexgen <board_size> <num_games> <seed>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "shannon_node_switching_game.h"

// restated from the reference's util.cpp:17-19 (util.cpp itself needs CrazyAra's customuci.h)
int repeatable_random_choice(std::vector<int>& vec) { return vec[rand() % vec.size()]; }

static void dump(const Node_switching_game& game, int game_id, int move_num) {
  torch::Device device(torch::kCPU, 0);
  std::vector<torch::Tensor> t = game.convert_graph(device);
  torch::Tensor x = t[0].contiguous();
  torch::Tensor ei = t[1].contiguous();
  const int64_t n = x.size(0), E = ei.size(1);
  std::printf("P %d %d %lld %lld %d\n", game_id, move_num, (long long)n, (long long)E,
              game.onturn == game.maker_color ? 1 : 0);
  const float* xp = x.data_ptr<float>();
  for (int64_t i = 0; i < n; ++i) std::printf("%d ", (int)xp[i * 3 + 0]);
  std::printf("\n");
  const int64_t* ep = ei.data_ptr<int64_t>();
  for (int64_t e = 0; e < E; ++e) std::printf("%lld ", (long long)ep[e]);
  std::printf("\n");
  for (int64_t e = 0; e < E; ++e) std::printf("%lld ", (long long)ep[E + e]);
  std::printf("\n");
}

int main(int argc, char** argv) {
  if (argc < 4) { std::fprintf(stderr, "usage: hexgen size games seed\n"); return 2; }
  const int size = std::atoi(argv[1]), games = std::atoi(argv[2]), seed = std::atoi(argv[3]);
  srand(seed);
  for (int g = 0; g < games; ++g) {
    Node_switching_game game(size, false);
    if (g % 2 == 0) game.switch_onturn();
    dump(game, g, 0);
    int mv = 0;
    while (true) {
      int move = game.get_random_action();
      game.make_move(move, false, NOPLAYER, true);
      ++mv;
      if (game.who_won() != NOPLAYER) break;
      dump(game, g, mv);
    }
  }
  return 0;
}
