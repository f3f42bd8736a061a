map_name =  "maps/GDC3.txt"
-- Simulator starting location.
start_x = 13.016267
start_y = 21.534157
start_angle = 0.0

human_config_list = { "config/human_crowd_scenario_configs/example_scenario/human/human_config_0.lua" ,
"config/human_crowd_scenario_configs/example_scenario/human/human_config_1.lua" ,
"config/human_crowd_scenario_configs/example_scenario/human/human_config_2.lua" ,
"config/human_crowd_scenario_configs/example_scenario/human/human_config_3.lua"  }

