// Generator (run in the build container only): prints, as JSON, the constant tables the reference's `oak` Python module
// exposes (cpp/src/pyoak.cc:586-636) by compiling the reference's own headers in place. tools/gen_oak_constants.py turns the
// output into oak_b200/oak_constants.py. Not part of the product or of the tests.
#include <encode/battle/battle.h>
#include <encode/battle/policy.h>
#include <encode/build/tensorizer.h>
#include <libpkmn/data/strings.h>
#include <nn/default-hyperparameters.h>

#include <cstdio>
#include <string>

template <size_t N, size_t M>
void labels(const char* name, const std::array<std::array<char, M>, N>& data, bool extra_empty = false) {
  printf("\"%s\": [", name);
  for (size_t i = 0; i < N; i++) printf("%s\"%s\"", i ? ", " : "", data[i].data());
  if (extra_empty) printf(", \"\"");
  printf("],\n");
}

int main() {
  using T = Encode::Build::Tensorizer<>;
  printf("{\n");
  printf("\"pokemon_in_dim\": %d, \"active_in_dim\": %d, \"policy_out_dim\": %d,\n", (int)Encode::Battle::Pokemon::n_dim,
         (int)Encode::Battle::ActivePokemon::n_dim, (int)Encode::Battle::Policy::n_dim);
  printf("\"pokemon_hidden_dim\": %d, \"pokemon_out_dim\": %d, \"active_hidden_dim\": %d, \"active_out_dim\": %d, \"side_out_dim\": %d, "
         "\"hidden_dim\": %d, \"value_hidden_dim\": %d, \"policy_hidden_dim\": %d,\n",
         (int)NN::Battle::Default::pokemon_hidden_dim, (int)NN::Battle::Default::pokemon_out_dim, (int)NN::Battle::Default::active_hidden_dim,
         (int)NN::Battle::Default::active_out_dim, (int)NN::Battle::Default::side_out_dim, (int)NN::Battle::Default::hidden_dim,
         (int)NN::Battle::Default::value_hidden_dim, (int)NN::Battle::Default::policy_hidden_dim);
  printf("\"build_policy_hidden_dim\": %d, \"build_value_hidden_dim\": %d, \"build_max_actions\": %d,\n",
         (int)NN::Build::Default::policy_hidden_dim, (int)NN::Build::Default::value_hidden_dim, (int)T::max_actions);
  printf("\"species_move_list\": [");
  for (int i = 0; i < (int)T::species_move_list_size; i++) {
    const auto p = T::species_move_list(i);
    printf("%s[%d, %d]", i ? ", " : "", (int)p.first, (int)p.second);
  }
  printf("],\n\"species_move_table\": [");
  for (size_t s = 0; s < T::SPECIES_MOVE_TABLE.size(); s++) {
    printf("%s[", s ? ", " : "");
    for (size_t m = 0; m < T::SPECIES_MOVE_TABLE[s].size(); m++) printf("%s%d", m ? ", " : "", T::SPECIES_MOVE_TABLE[s][m]);
    printf("]");
  }
  printf("],\n");
  labels("move_names", PKMN::Data::MOVE_CHAR_ARRAY);
  labels("species_names", PKMN::Data::SPECIES_CHAR_ARRAY);
  labels("active_dim_labels", Encode::Battle::Active::dim_labels);
  labels("pokemon_dim_labels", Encode::Battle::Pokemon::dim_labels);
  labels("active_pokemon_dim_labels", Encode::Battle::ActivePokemon::dim_labels);
  labels("policy_dim_labels", Encode::Battle::Policy::dim_labels, true);
  printf("\"source\": \"cpp/src/pyoak.cc:586-636\"\n}\n");
  return 0;
}
