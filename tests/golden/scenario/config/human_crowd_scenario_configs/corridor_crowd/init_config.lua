map_name =  "maps/AHG2.txt"
-- Simulator starting location.
start_x = 4.25
start_y = -11
start_angle = 1.5707963

human_config_list = { "config/human_crowd_scenario_configs/corridor_crowd/human/human_config_0.lua" ,
"config/human_crowd_scenario_configs/corridor_crowd/human/human_config_1.lua" ,
"config/human_crowd_scenario_configs/corridor_crowd/human/human_config_2.lua" ,
"config/human_crowd_scenario_configs/corridor_crowd/human/human_config_3.lua" ,
"config/human_crowd_scenario_configs/corridor_crowd/human/human_config_4.lua" ,
"config/human_crowd_scenario_configs/corridor_crowd/human/human_config_5.lua"  }

