int orc_euler_placeholder(void){return 0;}
